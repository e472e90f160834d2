"""N4: genuine ``.keras`` archives (zip: config.json + model.weights.h5) read without h5py / TensorFlow.
PARITY UNPINNED -- no genuine archive can be produced in this image; the HDF5 files come from tests/h5_writer.py, an
independent writer that follows the HDF5 file-format specification in two dialects."""
import importlib
import io
import json
import zipfile

import numpy as np
import pytest

import h5_writer
from conftest import PKG_NAME


def _config(uqk, names=None, rate2=None):
    layers, names = [{"class_name": "InputLayer", "config": {"name": "input_1", "batch_input_shape": [None, 60, 4]}}], list(names or [])
    for i, b in enumerate(uqk.arch.BLOCKS):
        nm = lambda kind: names.pop(0) if names else f"{kind}_{i + 1}"
        layers.append({"class_name": "Conv1D", "config": {"name": nm("conv"), "filters": b.cout, "kernel_size": [b.k], "strides": [1],
                                                          "padding": "same", "dilation_rate": [1], "activation": "relu", "use_bias": True}})
        layers.append({"class_name": "BatchNormalization", "config": {"name": nm("bn"), "axis": [2], "momentum": 0.99, "epsilon": 0.001,
                                                                      "center": True, "scale": True}})
        layers.append({"class_name": "Dropout", "config": {"name": nm("drop"), "rate": rate2 if (rate2 and i == 2) else b.rate, "noise_shape": None}})
    layers.append({"class_name": "GlobalAveragePooling1D", "config": {"name": "gap"}})
    layers.append({"class_name": "Dense", "config": {"name": "out", "units": 1, "activation": "sigmoid", "use_bias": True}})
    return {"class_name": "Sequential", "config": {"name": "Alarcon_1D_CNN_Model", "layers": layers}}


def _tree(uqk, ws, keras3):
    """model.weights.h5 of the architecture: <prefix>/<layer key>/vars/<i>."""
    layers = {}
    wi = 0
    counts = {"conv1d": 0, "batch_normalization": 0, "dense": 0}
    for i, b in enumerate(uqk.arch.BLOCKS):
        for kind, n, cfgname in (("conv1d", 2, f"conv_{i + 1}"), ("batch_normalization", 4, f"bn_{i + 1}")):
            k = counts[kind]; counts[kind] += 1
            key = (kind + (f"_{k}" if k else "")) if keras3 else cfgname
            layers[key] = {"vars": {str(j): ws[wi + j] for j in range(n)}}
            wi += n
    layers["dense" if keras3 else "out"] = {"vars": {"0": ws[36], "1": ws[37]}}
    top = "layers" if keras3 else "_layer_checkpoint_dependencies"
    return {top: layers, "optimizer": {"vars": {"0": np.zeros(3, np.float32)}}}


def _archive(cfg, h5_bytes):
    buf = io.BytesIO()
    with zipfile.ZipFile(buf, "w") as z:
        z.writestr("metadata.json", json.dumps({"keras_version": "2.12.0"}))
        z.writestr("config.json", json.dumps(cfg))
        z.writestr("model.weights.h5", h5_bytes)
    return buf.getvalue()


@pytest.mark.parametrize("dialect", ["v0_contiguous", "v0_chunked", "v2_compact"])
@pytest.mark.parametrize("keras3", [False, True])
def test_load_model_reads_a_keras_archive(tmp_path, uqk, dialect, keras3):
    model = importlib.import_module(PKG_NAME + ".model")
    ws = uqk.synthetic.make_member_weights(1)
    tree = _tree(uqk, ws, keras3)
    h5 = {"v0_contiguous": lambda: h5_writer.write_v0(tree), "v0_chunked": lambda: h5_writer.write_v0(tree, chunks="auto"),
          "v2_compact": lambda: h5_writer.write_v2(tree)}[dialect]
    if dialect == "v0_chunked":
        # one chunk shape per rank: ragged edges on every axis
        h5 = lambda: _chunked(tree)
    path = tmp_path / "AlCNN1D_no_pool.keras"
    path.write_bytes(_archive(_config(uqk), h5()))
    m = model.load_model(str(path))
    got = m.get_weights()
    assert len(got) == 38
    for a, b in zip(got, ws):
        np.testing.assert_array_equal(a, b)


def _chunked(tree):
    def conv(t):
        return {k: conv(v) if isinstance(v, dict) else v for k, v in t.items()}
    # write every dataset with chunks of (5, 3, 7)[:rank] -> partial edge chunks
    import h5_writer as hw
    orig = hw._dataset_v1

    def patched(buf, arr, chunks=None):
        return orig(buf, arr, tuple((5, 3, 7)[:arr.ndim]))
    hw._dataset_v1 = patched
    try:
        return hw.write_v0(conv(tree))
    finally:
        hw._dataset_v1 = orig


def test_reader_details_and_errors(tmp_path, uqk):
    ka = importlib.import_module(PKG_NAME + ".keras_archive")
    model = importlib.import_module(PKG_NAME + ".model")
    rng = np.random.default_rng(0)
    tree = {"a": {"b": {"vars": {"0": rng.standard_normal((3, 4)).astype(np.float32), "1": rng.standard_normal(5)}}},
            "many": {f"d{i:02d}": np.full((2,), i, np.float32) for i in range(21)}}      # 21 entries: three symbol-table nodes
    for blob in (h5_writer.write_v0(tree), h5_writer.write_v2(tree)):
        ds = ka._H5(blob).datasets()
        assert set(ds) == {"/a/b/vars/0", "/a/b/vars/1"} | {f"/many/d{i:02d}" for i in range(21)}
        np.testing.assert_array_equal(ds["/a/b/vars/0"], tree["a"]["b"]["vars"]["0"])
        assert ds["/a/b/vars/1"].dtype == np.float64
        np.testing.assert_array_equal(ds["/many/d20"], [20, 20])
    assert h5_writer.lookup3(b"") == 0xDEADBEEF and h5_writer.lookup3(b"Four score and seven years ago") == 0x17770551   # published vectors
    with pytest.raises(ka.H5FormatError):
        ka._H5(b"not hdf5 at all")
    ws = uqk.synthetic.make_member_weights(0)
    # config.json goes through validate_keras_model: a different dropout rate is refused before the weights are used
    bad = tmp_path / "bad.keras"
    bad.write_bytes(_archive(_config(uqk, rate2=0.25), h5_writer.write_v0(_tree(uqk, ws, False))))
    with pytest.raises(ValueError, match="Dropout rate"):
        model.load_model(str(bad))
    # a layer without variables in the file is reported by name
    t = _tree(uqk, ws, False)
    del t["_layer_checkpoint_dependencies"]["bn_3"]
    miss = tmp_path / "miss.keras"
    miss.write_bytes(_archive(_config(uqk), h5_writer.write_v0(t)))
    with pytest.raises(ValueError, match="bn_3"):
        model.load_model(str(miss))
