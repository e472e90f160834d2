#!/usr/bin/env python
"""Generates tests/golden/trained_member0.keras: a PSEUDO-TRAINED member of the 1D-CNN.

TEST INFRASTRUCTURE.  The Glorot-initialised synthetic members need a dense head scaled ~100x before their
probabilities leave 0.5, so bf16 rounding is amplified by the same factor and label checks on them say little.
This script trains the architecture of reference models/cnn_baseline_train.py:55-94 the way the reference does
(Adam, binary cross-entropy, BatchNormalization in training mode with momentum 0.99, Dropout active;
cnn_baseline_train.py:96-104, :195-216) for a few hundred steps of torch-CPU autograd on synthetic windows with a
learnable synthetic label, so that the deterministic probabilities spread over (0, 1).

    python tests/golden/make_trained_fixture.py [--steps 500] [--out tests/golden/trained_member0.keras]

Also writes trained_member0_golden.npz: float64-oracle probabilities of the trained member on 256 fixed windows
(deterministic and one MC-Dropout pass with Philox twin masks), used by the CPU tests to pin the fixture.
"""
import argparse
import importlib
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"


def synthetic_labels(x: np.ndarray, seed: int = 2026):
    """A label a small CNN can learn from standardised windows: late-vs-early level of channel 0 (a desaturation-like
    drift) plus the roughness of channel 2, through a noisy logistic link (so the Bayes-optimal p spreads over (0,1))."""
    drift = x[:, 40:, 0].mean(axis=1) - x[:, :20, 0].mean(axis=1)
    rough = np.abs(np.diff(x[:, :, 2], axis=1)).mean(axis=1)
    f = 1.2 * drift + 4.0 * (rough - np.median(rough))
    f = (f - f.mean()) / f.std()
    p = 1.0 / (1.0 + np.exp(-(2.0 * f - 0.8)))
    y = (np.random.default_rng(seed).random(len(x)) < p).astype(np.float32)
    return y, p


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--train-windows", type=int, default=16384)
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "trained_member0.keras"))
    ap.add_argument("--golden-only", action="store_true", help="keep the weights in --out, regenerate only the golden npz")
    args = ap.parse_args()
    pkg = importlib.import_module(PKG)
    model_mod = importlib.import_module(PKG + ".model")
    arch, syn = pkg.arch, pkg.synthetic
    torch.manual_seed(2025)
    rng = np.random.default_rng(2025)

    if args.golden_only:
        write_golden(args.out, model_mod.load_model(args.out).get_weights(), arch, syn)
        return

    x = syn.make_windows(args.train_windows, seed=777)
    y, _ = synthetic_labels(x)
    xt = torch.from_numpy(x.astype(np.float32)).permute(0, 2, 1).contiguous()
    yt = torch.from_numpy(y)

    # Keras initialisation: Glorot-uniform kernels, zero biases, gamma 1, beta 0, moving mean 0 / variance 1
    params, moving = [], []
    for b in arch.BLOCKS:
        lim = np.sqrt(6.0 / (b.k * b.cin + b.k * b.cout))
        w = torch.tensor(rng.uniform(-lim, lim, size=(b.cout, b.cin, b.k)).astype(np.float32), requires_grad=True)
        params += [w, torch.zeros(b.cout, requires_grad=True), torch.ones(b.cout, requires_grad=True),
                   torch.zeros(b.cout, requires_grad=True)]
        moving += [torch.zeros(b.cout), torch.ones(b.cout)]
    lim = np.sqrt(6.0 / (arch.DENSE_IN + 1))
    wd = torch.tensor(rng.uniform(-lim, lim, size=(arch.DENSE_IN,)).astype(np.float32), requires_grad=True)
    bd = torch.zeros(1, requires_grad=True)
    params += [wd, bd]
    opt = torch.optim.Adam(params, lr=1e-3, eps=1e-7)          # Keras Adam defaults

    def forward_train(a):
        for i, b in enumerate(arch.BLOCKS):
            w, bi, g, be = params[4 * i:4 * i + 4]
            a = F.relu(F.conv1d(a, w, bi, padding=b.halo))
            mu = a.mean(dim=(0, 2))
            var = a.var(dim=(0, 2), unbiased=False)
            with torch.no_grad():                               # Keras: moving = moving * 0.99 + batch * 0.01
                moving[2 * i].mul_(arch.BN_MOMENTUM).add_(mu, alpha=1 - arch.BN_MOMENTUM)
                moving[2 * i + 1].mul_(arch.BN_MOMENTUM).add_(var, alpha=1 - arch.BN_MOMENTUM)
            a = (a - mu[None, :, None]) * (g / torch.sqrt(var + arch.BN_EPS))[None, :, None] + be[None, :, None]
            a = F.dropout(a, p=b.rate, training=True)
        return a.mean(dim=2) @ wd + bd

    n = xt.shape[0]
    for step in range(args.steps):
        idx = torch.from_numpy(rng.integers(0, n, size=args.batch))
        loss = F.binary_cross_entropy_with_logits(forward_train(xt[idx]), yt[idx])
        opt.zero_grad()
        loss.backward()
        opt.step()
        if step % 50 == 0 or step == args.steps - 1:
            print(f"step {step:4d}  loss {float(loss):.4f}", flush=True)

    ws = []
    for i, b in enumerate(arch.BLOCKS):
        w, bi, g, be = [t.detach().numpy() for t in params[4 * i:4 * i + 4]]
        ws += [np.ascontiguousarray(w.transpose(2, 1, 0)), bi, g, be, moving[2 * i].numpy(), moving[2 * i + 1].numpy()]
    ws += [wd.detach().numpy().reshape(arch.DENSE_IN, 1), bd.detach().numpy()]
    ws = [np.asarray(w, dtype=np.float32) for w in ws]
    model_mod.save_model(model_mod.UQModel(ws), args.out)
    write_golden(args.out, ws, arch, syn)


def write_golden(out, ws, arch, syn):
    from oracle import network, philox
    xg = syn.make_windows(256, seed=4242)
    det = network.forward_f64(ws, xg)[:, 0]
    masks = [philox.dropout_keep_mask(2025, np.arange(256), 0, i, b.cout, b.rate) for i, b in enumerate(arch.BLOCKS)]
    mcd = network.forward_f64(ws, xg, training=True, masks=masks)[:, 0]
    yg, _ = synthetic_labels(xg)
    np.savez(out.replace(".keras", "_golden.npz"), det=det, mcd_pass0_seed2025=mcd, y=yg)
    q = np.quantile(det, [0.01, 0.1, 0.5, 0.9, 0.99])
    print("deterministic p quantiles (1/10/50/90/99 %):", np.round(q, 4), " accuracy vs synthetic label:",
          float(((det > 0.5) == (yg > 0.5)).mean()), " fraction |p-0.5|<0.01:", float((np.abs(det - 0.5) < 0.01).mean()))


if __name__ == "__main__":
    main()
