#!/usr/bin/env python
"""Opcode mix and warp-stall breakdown of one kernel from `ncu -i rep --page source --csv --print-source sass`.
usage: ncu -i x.ncu-rep --page source --csv --print-source sass > src.csv; tools/ncu_source_summary.py src.csv out.txt [note]
Instructions are split into "work" and "waiting" (mbarrier polls, sleeps, named barriers and their loop control)."""
import collections, csv, sys

WAIT_OPS = ("SYNCS", "NANOSLEEP", "BSSY", "BSYNC", "NOP", "BAR", "WARPSYNC", "UMOV", "BRA")


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hi = [i for i, r in enumerate(rows) if "Source" in r and "Address" in r][0]
    hdr, data = rows[hi], rows[hi + 1:]
    isrc, iinst, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    inst, samp, st = collections.Counter(), collections.Counter(), collections.Counter()
    for r in data:
        t = r[isrc].split()
        if not t:
            continue
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        inst[op] += int(r[iinst] or 0)
        samp[op] += int(r[isamp] or 0)
        for h in stalls:
            st[(op in WAIT_OPS, h)] += int(r[hdr.index(h)] or 0)
    tot = sum(inst.values())
    out = [f"# {sys.argv[3] if len(sys.argv) > 3 else ''}", f"warp instructions executed: {tot}", "",
           "opcode        share of instructions   samples"]
    for op, c in inst.most_common(22):
        out.append(f"{op:12s} {100 * c / tot:6.2f} %              {samp[op]}")
    for flag, name in ((False, "work instructions"), (True, "waiting (mbarrier polls, sleeps, named barriers, loop control)")):
        s = sum(v for (f, _), v in st.items() if f == flag)
        out += ["", f"stall samples on {name}: {s}"]
        for (f, h), v in sorted(st.items(), key=lambda kv: -kv[1]):
            if f == flag and v > 0.03 * s:
                out.append(f"   {h:26s} {100 * v / s:5.1f} %")
    open(sys.argv[2], "w").write("\n".join(out) + "\n")
    print("\n".join(out))


if __name__ == "__main__":
    main()
