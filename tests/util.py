"""Shared helpers for the parity tests (seeded synthetic inputs of the BASELINE.json shapes)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def hashgrid_config(bound=2.0, table_std=0.5, seed=0, level_dim=2, num_levels=16, log2_hashmap_size=19):
    """Config-2 encoder: 16 levels x 2 feats, 2^19 rows cap, base 16 -> 2048*bound (network.py:31, encoding.py:45-49)."""
    from oracle import oracle as O
    offsets, pls = O.level_table(3, num_levels, level_dim, 2.0, 16, log2_hashmap_size, 2048 * bound)
    rng = np.random.default_rng(seed)
    table = rng.uniform(-table_std, table_std, (int(offsets[-1]), level_dim)).astype(np.float32)
    return dict(offsets=offsets, per_level_scale=pls, S=float(np.log2(pls)), H=16, table=table, bound=bound,
                L=num_levels, C=level_dim)


def sigma_net_weights(seed=1, in_dim=32, hidden=64, out=16):
    """torch nn.Linear default init: U(-1/sqrt(in), 1/sqrt(in)), bias-free (network.py:33-47)."""
    rng = np.random.default_rng(seed)
    W0 = rng.uniform(-1, 1, (hidden, in_dim)).astype(np.float32) / np.sqrt(in_dim).astype(np.float32)
    W1 = rng.uniform(-1, 1, (out, hidden)).astype(np.float32) / np.sqrt(hidden).astype(np.float32)
    return W0, W1


def rel_err(a, b, floor=1e-30):
    """max |a-b| / max|b| — the relative error used for the fp32 (1e-5) and fp16 (2e-3) bars."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), floor))


def load_ref(name):
    """Reference CUDA extension built by oracle/build_ref.py, or None."""
    from oracle import build_ref
    return build_ref.load_ref(name)
