"""Fixed architecture of the 1D-CNN whose forward passes are the hot path.

Mirrors ``al_1d_cnn_create_model`` (reference models/cnn_baseline_train.py:55-94, identical copy
models/train_deep_ensemble_cnns.py:30-71): six blocks of
Conv1D(same, ReLU) -> BatchNormalization -> Dropout, then GlobalAveragePooling1D and Dense(1, sigmoid).
Input windows are (60, 4) (reference cnn_baseline_train.py:169-170).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

WINDOW_LEN = 60          # time steps per window
WINDOW_FEATURES = 4      # SaO2, PR, THOR RES, ABDO RES (prepare_numpy_datasets.py:53)
BN_EPS = 1e-3            # Keras BatchNormalization default epsilon
BN_MOMENTUM = 0.99       # Keras BatchNormalization default momentum
N_BLOCKS = 6


@dataclass(frozen=True)
class Block:
    cin: int
    cout: int
    k: int
    rate: float  # dropout rate (probability of zeroing)

    @property
    def halo(self) -> int:
        return (self.k - 1) // 2


# (Cin, Cout, kernel, dropout rate) -- cnn_baseline_train.py:59-86
BLOCKS: Tuple[Block, ...] = (
    Block(4, 128, 7, 0.3),
    Block(128, 192, 5, 0.3),
    Block(192, 224, 3, 0.4),
    Block(224, 96, 7, 0.2),
    Block(96, 256, 9, 0.3),
    Block(256, 96, 9, 0.5),
)
DENSE_IN = 96
MAX_CH = 256

# Number of float32 values in one member's weight blob, Keras get_weights() order:
# per block [kernel (k,Cin,Cout), bias, gamma, beta, moving_mean, moving_var], then dense kernel (96,1), bias (1,)
def blob_layout():
    """Returns list of (name, shape) in Keras ``get_weights()`` order (38 arrays)."""
    out = []
    for i, b in enumerate(BLOCKS):
        out.append((f"conv{i+1}/kernel", (b.k, b.cin, b.cout)))
        out.append((f"conv{i+1}/bias", (b.cout,)))
        out.append((f"bn{i+1}/gamma", (b.cout,)))
        out.append((f"bn{i+1}/beta", (b.cout,)))
        out.append((f"bn{i+1}/moving_mean", (b.cout,)))
        out.append((f"bn{i+1}/moving_variance", (b.cout,)))
    out.append(("dense/kernel", (DENSE_IN, 1)))
    out.append(("dense/bias", (1,)))
    return out


def blob_size() -> int:
    n = 0
    for _, shp in blob_layout():
        k = 1
        for s in shp:
            k *= s
        n += k
    return n


# Algorithmic FLOP per window-pass: dense count (padded taps counted) and valid-tap count.
def flops_per_window_pass(valid_taps: bool = True) -> int:
    total = 0
    for b in BLOCKS:
        taps = WINDOW_LEN * b.k
        if valid_taps:
            taps -= b.halo * (b.halo + 1)
        total += 2 * taps * b.cin * b.cout
    total += 2 * DENSE_IN
    return total


def dropout_threshold_u16(rate: float) -> int:
    """16-bit integer threshold: an element is KEPT iff its 16-bit uniform draw >= threshold
    (TF semantics: keep iff u >= rate, u in [0,1)); csrc/common.cuh ``drop_thr16``."""
    return int(round(rate * 65536.0))
