"""Oracle network restatement: self-consistency (float64 vs float32 torch), regression against the
committed oracle outputs, Philox known-answer vectors (Random123) and mask statistics."""
import os

import numpy as np

from oracle import network, philox


def test_philox_random123_kat():
    # Random123 kat_vectors, philox4x32-10
    cases = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
             ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
             ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
              (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for c, k, want in cases:
        out = philox.philox4x32(*[np.array([x]) for x in c], k[0], k[1])
        assert tuple(int(o[0]) for o in out) == want


def test_mask_keep_rates(uqk):
    for i, b in enumerate(uqk.arch.BLOCKS):
        m = philox.dropout_keep_mask(99, np.arange(200), 1, i, b.cout, b.rate)
        n = m.size
        sigma = np.sqrt(b.rate * (1 - b.rate) / n)
        assert abs(m.mean() - (1 - b.rate)) < 4 * sigma + 2e-5
    w = philox.pack_mask_bits(m)
    assert w.shape == (200, 60, 8) and w.dtype == np.uint32
    assert ((w[..., 0] >> 5) & 1).astype(bool).tolist() == m[..., 5].tolist()


def test_forward_regression_and_f32(uqk, golden_dir):
    g = np.load(os.path.join(golden_dir, "forward_oracle_golden.npz"))
    X = uqk.synthetic.make_windows(24, seed=11)
    ws = uqk.synthetic.make_member_weights(0)
    p = network.forward_f64(ws, X)
    np.testing.assert_allclose(p, g["det_member0"], rtol=1e-9, atol=1e-12)
    p32 = network.forward_f32_torch(ws, X)
    assert np.abs(p32 - p).max() < 1e-5         # fp32 vs fp64 tolerance of the high-precision mode
    p32b = network.forward_f32_torch(ws, X, batch_size=7)
    assert np.abs(p32b - p).max() < 1e-5
    masks = [philox.dropout_keep_mask(2025, np.arange(24), 1, i, b.cout, b.rate)
             for i, b in enumerate(uqk.arch.BLOCKS)]
    pt = network.forward_f64(ws, X, training=True, masks=masks)
    np.testing.assert_allclose(pt, g["mcd_batch_t1"], rtol=1e-9, atol=1e-12)
    pt32 = network.forward_f32_torch(ws, X, training=True, masks=masks)
    assert np.abs(pt32 - pt).max() < 1e-4


def test_torch_f64_restatement_equals_numpy_f64(uqk):
    """forward_f64_torch (used by the T=50 parity tests) is the same function as forward_f64."""
    X = uqk.synthetic.make_windows(40, seed=13)
    ws = uqk.synthetic.make_member_weights(1)
    masks = [philox.dropout_keep_mask(5, np.arange(40), 2, i, b.cout, b.rate) for i, b in enumerate(uqk.arch.BLOCKS)]
    _, st = network.forward_f64(ws, X, training=True, masks=masks, return_stats=True)
    for kw in (dict(), dict(training=True, masks=masks), dict(training=True, masks=masks, bn_mode="inference"),
               dict(training=True, masks=masks, bn_stats=st)):
        np.testing.assert_allclose(network.forward_f64_torch(ws, X, **kw), network.forward_f64(ws, X, **kw), rtol=0, atol=1e-12)


def test_trained_fixture_is_pinned_to_its_golden(uqk, golden_dir):
    """tests/golden/trained_member0.keras (made by make_trained_fixture.py) loads through the package's reader and the
    oracle reproduces the committed outputs; its probabilities spread over (0, 1)."""
    import importlib
    from conftest import PKG_NAME
    model = importlib.import_module(PKG_NAME + ".model")
    ws = model.load_model(os.path.join(golden_dir, "trained_member0.keras")).get_weights()
    g = np.load(os.path.join(golden_dir, "trained_member0_golden.npz"))
    X = uqk.synthetic.make_windows(256, seed=4242)
    det = network.forward_f64(ws, X)[:, 0]
    np.testing.assert_allclose(det, g["det"], rtol=0, atol=1e-12)
    masks = [philox.dropout_keep_mask(2025, np.arange(256), 0, i, b.cout, b.rate) for i, b in enumerate(uqk.arch.BLOCKS)]
    np.testing.assert_allclose(network.forward_f64(ws, X, training=True, masks=masks)[:, 0], g["mcd_pass0_seed2025"], rtol=0, atol=1e-12)
    assert np.quantile(det, 0.1) < 0.1 and np.quantile(det, 0.9) > 0.8


def test_flop_counts(uqk):
    # SURVEY.md section 8(d)
    assert uqk.arch.flops_per_window_pass(False) == 101_806_272
    assert uqk.arch.flops_per_window_pass(True) == 98_844_864
    assert uqk.arch.blob_size() == 853_441


def test_standardize_matches_reference_golden(uqk, golden_dir):
    g = np.load(os.path.join(golden_dir, "standardize_golden.npz"))
    np.testing.assert_allclose(uqk.synthetic.standardize_per_window(g["raw"]), g["out"], rtol=1e-13, atol=1e-13)
