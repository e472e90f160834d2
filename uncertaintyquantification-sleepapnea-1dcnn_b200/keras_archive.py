"""Reader for genuine Keras ``.keras`` archives.

The reference saves its models with ``model.save("*.keras")`` (models/cnn_baseline_train.py:230,
models/train_deep_ensemble_cnns.py:170) and the drivers load those files
(analyze_mcd_patient_level.py:199, analyze_de_patient_level.py:40-58).  A ``.keras`` file is a zip holding
``config.json`` (architecture) and ``model.weights.h5`` (HDF5).  TensorFlow and h5py are absent from this image, so the
HDF5 part is read by the minimal pure-Python parser below (h5py is used instead when it can be imported):

* superblock versions 0 / 1 (what h5py writes by default) and 2 / 3;
* groups: symbol tables (v1 B-tree + local heap + SNOD nodes) and compact link messages of version-2 object headers
  (dense link storage in a fractal heap is NOT supported -- it appears only in groups with more than 8 links created
  with ``libver='latest'``);
* datasets: little-endian IEEE float32 / float64, simple dataspaces, data layout version 3: compact, contiguous and
  chunked (v1 B-tree) WITHOUT filters (Keras writes its variables uncompressed and unchunked).

PARITY UNPINNED: no genuine archive exists in this image and nothing here can produce one; the parser is tested against
files written by an independent minimal writer that follows the HDF5 file-format specification (tests/h5_writer.py).
A user with TensorFlow can confirm it with
``np.testing.assert_equal(load_model(p).get_weights(), tf.keras.models.load_model(p).get_weights())``.
"""
from __future__ import annotations

import io
import json
import re
import struct
import zipfile
from typing import Dict, List, Tuple

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"


class H5FormatError(ValueError):
    pass


class _H5:
    """Minimal HDF5 reader over a bytes object: ``datasets()`` -> {"/group/.../name": ndarray}."""

    def __init__(self, data: bytes):
        self.d = data
        base = data.find(SIGNATURE)
        if base != 0:
            raise H5FormatError("not an HDF5 file (signature missing at offset 0)")
        ver = data[8]
        if ver in (0, 1):
            self.O, self.L = data[13], data[14]
            p = 24 + (4 if ver == 1 else 0)
            p += 4 * self.O                                  # base, free-space, end-of-file, driver-info addresses
            _, self.root_header, cache = self._symbol_entry(p)[:3]
        elif ver in (2, 3):
            self.O, self.L = data[9], data[10]
            p = 12 + 3 * self.O                              # base, superblock extension, end-of-file addresses
            self.root_header = self._addr(p)
        else:
            raise H5FormatError(f"unsupported HDF5 superblock version {ver}")
        if self.O != 8 or self.L != 8:
            raise H5FormatError(f"unsupported offset/length sizes {self.O}/{self.L} (need 8/8)")

    # ---- primitives ----------------------------------------------------------------------------------------------------
    def _u(self, p, n):
        return int.from_bytes(self.d[p:p + n], "little")

    def _addr(self, p):
        return self._u(p, self.O)

    def _symbol_entry(self, p):
        name_off, header, cache = self._addr(p), self._addr(p + self.O), self._u(p + 2 * self.O, 4)
        scratch = p + 2 * self.O + 8
        return name_off, header, cache, scratch

    # ---- object headers --------------------------------------------------------------------------------------------------
    def _messages(self, addr) -> List[Tuple[int, bytes]]:
        """All header messages (type, body) of the object at ``addr``, continuation blocks followed."""
        d, out = self.d, []
        if d[addr:addr + 4] == b"OHDR":                      # version 2
            if d[addr + 4] != 2:
                raise H5FormatError("unsupported object header version")
            flags = d[addr + 5]
            p = addr + 6
            if flags & 0x20:
                p += 16                                      # four timestamps
            if flags & 0x10:
                p += 4                                       # max compact / min dense attributes
            szlen = 1 << (flags & 3)
            chunk0 = self._u(p, szlen)
            p += szlen
            blocks = [(p, chunk0)]
            track = bool(flags & 0x04)
            while blocks:
                p, size = blocks.pop(0)
                end = p + size
                while p + 4 <= end:
                    mtype, msize = d[p], self._u(p + 1, 2)
                    p += 4 + (2 if track else 0)
                    body = d[p:p + msize]
                    p += msize
                    if mtype == 0x10:
                        caddr, clen = self._addr_b(body, 0), self._u_b(body, self.O, self.L)
                        if d[caddr:caddr + 4] != b"OCHK":
                            raise H5FormatError("bad object header continuation")
                        blocks.append((caddr + 4, clen - 8))  # signature in front, checksum behind
                    elif mtype != 0:
                        out.append((mtype, body))
            return out
        if d[addr] != 1:
            raise H5FormatError(f"unsupported object header at {addr}")
        nmsg, size = self._u(addr + 2, 2), self._u(addr + 8, 4)
        blocks = [(addr + 16, size)]
        while blocks and len(out) < nmsg + 64:
            p, size = blocks.pop(0)
            end = p + size
            while p + 8 <= end:
                mtype, msize = self._u(p, 2), self._u(p + 2, 2)
                body = d[p + 8:p + 8 + msize]
                p += 8 + msize
                if mtype == 0x10:
                    blocks.append((self._addr_b(body, 0), self._u_b(body, self.O, self.L)))
                elif mtype != 0:
                    out.append((mtype, body))
        return out

    @staticmethod
    def _u_b(b, p, n):
        return int.from_bytes(b[p:p + n], "little")

    def _addr_b(self, b, p):
        return int.from_bytes(b[p:p + self.O], "little")

    # ---- groups ----------------------------------------------------------------------------------------------------------
    def _heap_name(self, heap_addr, off):
        d = self.d
        if d[heap_addr:heap_addr + 4] != b"HEAP":
            raise H5FormatError("bad local heap")
        seg = self._addr(heap_addr + 8 + 2 * self.L)
        end = d.index(b"\0", seg + off)
        return d[seg + off:end].decode("utf-8")

    def _btree_group(self, addr, heap, out):
        d = self.d
        if d[addr:addr + 4] != b"TREE" or d[addr + 4] != 0:
            raise H5FormatError("bad group B-tree node")
        level, used = d[addr + 5], self._u(addr + 6, 2)
        p = addr + 8 + 2 * self.O
        for i in range(used):
            child = self._addr(p + self.L)                   # key_i (L bytes), child_i (O bytes)
            p += self.L + self.O
            if level > 0:
                self._btree_group(child, heap, out)
                continue
            if d[child:child + 4] != b"SNOD":
                raise H5FormatError("bad symbol table node")
            q = child + 8
            for _ in range(self._u(child + 6, 2)):
                name_off, header = self._symbol_entry(q)[:2]
                out.append((self._heap_name(heap, name_off), header))
                q += 2 * self.O + 24

    def _links(self, addr) -> List[Tuple[str, int]]:
        out: List[Tuple[str, int]] = []
        for mtype, body in self._messages(addr):
            if mtype == 0x11:                                # symbol table message: old-style group
                self._btree_group(self._addr_b(body, 0), self._addr_b(body, self.O), out)
            elif mtype == 0x06:                              # link message (compact new-style group)
                flags, p = body[1], 2
                ltype = 0
                if flags & 0x08:
                    ltype = body[p]; p += 1
                if flags & 0x04:
                    p += 8
                if flags & 0x10:
                    p += 1
                nlen = 1 << (flags & 3)
                n = self._u_b(body, p, nlen); p += nlen
                name = body[p:p + n].decode("utf-8"); p += n
                if ltype == 0:
                    out.append((name, self._addr_b(body, p)))
            elif mtype == 0x02:                              # link info: dense storage?
                flags = body[1]
                p = 2 + (8 if flags & 1 else 0)
                if self._addr_b(body, p) != UNDEF:
                    raise H5FormatError("dense link storage (fractal heap) is not supported by the minimal HDF5 reader; "
                                        "install h5py or re-save the weights with np.savez")
        return out

    # ---- datasets --------------------------------------------------------------------------------------------------------
    def _dataset(self, msgs) -> np.ndarray:
        shape = dtype = layout = None
        for mtype, b in msgs:
            if mtype == 0x01:
                ver, rank, flags = b[0], b[1], b[2]
                p = 8 if ver == 1 else 4
                shape = tuple(self._u_b(b, p + 8 * i, 8) for i in range(rank))
            elif mtype == 0x03:
                cls, size = b[0] & 0x0F, self._u_b(b, 4, 4)
                if cls != 1 or (b[1] & 1) or size not in (4, 8):
                    raise H5FormatError("only little-endian float32 / float64 datasets are supported")
                dtype = np.dtype("<f4") if size == 4 else np.dtype("<f8")
            elif mtype == 0x0B:
                raise H5FormatError("filtered (compressed) datasets are not supported")
            elif mtype == 0x08:
                layout = b
        if shape is None or dtype is None or layout is None:
            raise H5FormatError("dataset header lacks dataspace / datatype / layout")
        count = int(np.prod(shape)) if shape else 1
        if layout[0] != 3:
            raise H5FormatError(f"unsupported data layout message version {layout[0]}")
        cls = layout[1]
        if cls == 0:                                         # compact
            size = self._u_b(layout, 2, 2)
            raw = layout[4:4 + size]
        elif cls == 1:                                       # contiguous
            addr, size = self._addr_b(layout, 2), self._u_b(layout, 2 + self.O, self.L)
            raw = b"" if addr == UNDEF else self.d[addr:addr + size]
            if addr == UNDEF:
                return np.zeros(shape, dtype=dtype)
        elif cls == 2:                                       # chunked, v1 B-tree
            ndim = layout[2]
            btree = self._addr_b(layout, 3)
            cdims = tuple(self._u_b(layout, 3 + self.O + 4 * i, 4) for i in range(ndim))[:-1]
            out = np.zeros(shape, dtype=dtype)
            if btree != UNDEF:
                self._chunks(btree, ndim, cdims, dtype, out)
            return out
        else:
            raise H5FormatError(f"unsupported layout class {cls}")
        return np.frombuffer(raw[:count * dtype.itemsize], dtype=dtype).reshape(shape).copy()

    def _chunks(self, addr, ndim, cdims, dtype, out):
        d = self.d
        if d[addr:addr + 4] != b"TREE" or d[addr + 4] != 1:
            raise H5FormatError("bad chunk B-tree node")
        level, used = d[addr + 5], self._u(addr + 6, 2)
        p = addr + 8 + 2 * self.O
        ksize = 8 + 8 * ndim
        for _ in range(used):
            csize, mask = self._u(p, 4), self._u(p + 4, 4)
            offs = tuple(self._u(p + 8 + 8 * i, 8) for i in range(ndim - 1))
            child = self._addr(p + ksize)
            p += ksize + self.O
            if level > 0:
                self._chunks(child, ndim, cdims, dtype, out)
                continue
            if mask:
                raise H5FormatError("filtered chunks are not supported")
            chunk = np.frombuffer(d[child:child + csize], dtype=dtype)[:int(np.prod(cdims))].reshape(cdims)
            sel = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, cdims, out.shape))
            out[sel] = chunk[tuple(slice(0, s.stop - s.start) for s in sel)]

    # ---- walk --------------------------------------------------------------------------------------------------------------
    def datasets(self) -> Dict[str, np.ndarray]:
        out: Dict[str, np.ndarray] = {}
        seen = set()

        def walk(addr, path):
            if addr in seen:
                return
            seen.add(addr)
            msgs = self._messages(addr)
            types = {t for t, _ in msgs}
            if 0x08 in types and 0x01 in types:
                out[path or "/"] = self._dataset(msgs)
                return
            for name, child in self._links(addr):
                walk(child, f"{path}/{name}")
        walk(self.root_header, "")
        return out


def read_h5_datasets(data: bytes) -> Dict[str, np.ndarray]:
    """{"/path/to/dataset": array} for every dataset of an HDF5 file image (h5py when importable, else the parser above)."""
    try:
        import h5py                                          # noqa: F401
    except Exception:
        return _H5(data).datasets()
    out: Dict[str, np.ndarray] = {}
    with h5py.File(io.BytesIO(data), "r") as f:
        f.visititems(lambda name, obj: out.__setitem__("/" + name, np.asarray(obj)) if isinstance(obj, h5py.Dataset) else None)
    return out


# ---- the .keras container ------------------------------------------------------------------------------------------------
class _ConfigLayer:
    """Duck-typed stand-in for a Keras layer built from config.json (what ``model.validate_keras_model`` inspects)."""

    def __init__(self, class_name: str, config: dict):
        self._config = dict(config)
        self.name = config.get("name", class_name)
        self.__class__ = type(class_name, (_ConfigLayer,), {})

    def get_config(self):
        return dict(self._config)


def _snake(name: str) -> str:
    """Keras' to_snake_case: Conv1D -> conv1d, BatchNormalization -> batch_normalization."""
    s = re.sub(r"(.)([A-Z][a-z0-9]+)", r"\1_\2", name)
    return re.sub(r"([a-z])([A-Z])", r"\1_\2", s).lower()


def read_keras_archive(path_or_bytes):
    """-> (layers from config.json as duck-typed objects, {layer key: [arrays in variable order]}).
    A layer's variables are the datasets ``<...>/<layer key>/vars/<i>`` of model.weights.h5 (Keras 2.12-2.15 and 3.x
    both use a ``vars`` group per layer; only the prefix in front of the layer key differs between the versions)."""
    data = path_or_bytes if isinstance(path_or_bytes, (bytes, bytearray)) else open(path_or_bytes, "rb").read()
    with zipfile.ZipFile(io.BytesIO(data)) as z:
        names = set(z.namelist())
        if "config.json" not in names or "model.weights.h5" not in names:
            raise ValueError("not a Keras v3 archive: config.json / model.weights.h5 missing")
        config = json.loads(z.read("config.json"))
        ds = read_h5_datasets(z.read("model.weights.h5"))
    cfg = config.get("config", config)
    layer_cfgs = cfg.get("layers", [])
    layers = [_ConfigLayer(lc.get("class_name", "?"), lc.get("config", {})) for lc in layer_cfgs]
    per_layer: Dict[str, Dict[int, np.ndarray]] = {}
    for p, arr in ds.items():
        parts = [q for q in re.split(r"[/\\]", p) if q]
        if len(parts) >= 3 and parts[-2] == "vars" and parts[-1].isdigit() and "optimizer" not in parts:
            per_layer.setdefault(parts[-3], {})[int(parts[-1])] = arr
    return layers, {k: [v[i] for i in sorted(v)] for k, v in per_layer.items()}


def weights_from_keras_archive(path_or_bytes) -> Tuple[list, List[np.ndarray]]:
    """-> (duck-typed layer list, the 38 arrays in ``get_weights()`` order).  Layers are matched to their variable
    groups by name (Keras 2.x) or by the snake-cased class name with an occurrence suffix (Keras 3: conv1d, conv1d_1, ...)."""
    layers, groups = read_keras_archive(path_or_bytes)
    weights: List[np.ndarray] = []
    seen: Dict[str, int] = {}
    for layer in layers:
        cls = type(layer).__name__
        k = seen.get(cls, 0)
        seen[cls] = k + 1
        if cls in ("InputLayer", "Dropout", "GlobalAveragePooling1D", "Activation", "Flatten"):
            continue                                         # no variables
        by_class = _snake(cls) + (f"_{k}" if k else "")
        key = layer.name if layer.name in groups else by_class if by_class in groups else None
        if key is None:
            raise ValueError(f"model.weights.h5 holds no variables for layer '{layer.name}' ({cls}); groups: {sorted(groups)}")
        weights += [np.asarray(a, dtype=np.float32) for a in groups[key]]
    return layers, weights
