"""Cluster signalling of the CTA-pair kernels and the failure path.

* Peer -> leader signals are relaxed cluster-scope mbarrier arrives by default ("CTA pairs"); variant bit 0
  switches to the release / acquire pairing the PTX memory model spells out.  10^4 small launches of every pair kernel,
  in both forms, are compared bit for bit with the kernel that uses no cluster at all (variant 16).
* The barrier watchdog reports through mapped host memory: a kernel that waits on a barrier nobody signals makes the
  next API call return UQK_E_KERNEL with the barrier code (run in a child process: the CUDA context is lost)."""
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import PKG_NAME, ROOT
from gpu_common import engine_with_members

pytestmark = pytest.mark.gpu


def test_pair_kernels_relaxed_and_release_acquire_match_the_single_cta_kernel(uqk):
    eng = engine_with_members()
    rng = np.random.default_rng(5)
    cases = []
    x_all = torch.from_numpy(uqk.synthetic.make_windows(96, seed=811)).cuda()
    for i in range(25):
        n, T, seed = int(rng.integers(1, 97)), int(rng.integers(1, 4)), int(rng.integers(1, 1 << 30))
        eng.option("variant", 16)
        ref = eng.mcd(x_all[:n], n_passes=T, seed=seed, bn_mode="inference", return_probs=True, metrics=(), device_out=True)["probs"].clone()
        cases.append((n, T, seed, ref))
    launches = 0
    try:
        for variant in (0, 1, 32, 33):               # two-tile pairs / one-tile pairs, relaxed / release-acquire
            eng.option("variant", variant)
            for rep in range(100):
                for n, T, seed, ref in cases:
                    got = eng.mcd(x_all[:n], n_passes=T, seed=seed, bn_mode="inference", return_probs=True, metrics=(), device_out=True)["probs"]
                    launches += 1
                    if rep % 10 == 0 or variant in (1, 33):
                        assert torch.equal(got, ref), (variant, n, T, seed)
        torch.cuda.synchronize()
        assert torch.equal(got, cases[-1][3])
        # the exact-mode sweeps (statistics kernel instantiation) in both forms
        x = x_all[:50]
        eng.option("variant", 0)
        a = eng.mcd(x, n_passes=3, seed=9, bn_mode="batch", precision="bf16", return_probs=True)["probs"]
        for pipelined in (1, 0):                      # pipelined sweeps (default) and the lock-step form
            eng.option("exact_pipelined", pipelined)
            eng.option("variant", 1)
            for _ in range(10):
                b = eng.mcd(x, n_passes=3, seed=9, bn_mode="batch", precision="bf16", return_probs=True)["probs"]
                np.testing.assert_array_equal(a, b)
    finally:
        eng.option("variant", 0)
        eng.option("exact_pipelined", 1)
    assert launches == 10_000
    eng.poll_error()


def test_watchdog_reports_through_host_memory_after_the_trap():
    code = f"""
import sys, importlib, ctypes as C
sys.path.insert(0, {ROOT!r})
lib = importlib.import_module({PKG_NAME!r} + '._lib')
engine = importlib.import_module({PKG_NAME!r} + '.engine')
eng = engine.Engine(0)
rc = eng.lib.uqk_debug_trip_watchdog(eng._h)
msg = eng.lib.uqk_last_error().decode()
assert rc == -7 and 'watchdog 99' in msg, (rc, msg)
assert eng.lib.uqk_poll_error(eng._h) == -7
try:
    eng.poll_error()
except lib.UqkError as e:
    assert 'code -7' in str(e) and 'restart' in str(e)
    print('WATCHDOG_OK')
"""
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert "WATCHDOG_OK" in r.stdout, (r.stdout[-500:], r.stderr[-1500:])
