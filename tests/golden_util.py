"""Seeded generators shared by tools/gen_golden_ref.py (which ran the reference on a B200) and the tests that
check the oracle / the CUDA path against tests/golden/ref_*.npz.  Changing anything here invalidates the fixtures."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

N_INFER = 4096              # multiple of 256: the reference pads n to 256 and touches those rows
CURVE_STEPS = 1000
GRID_SAMPLE_STRIDE = 61
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def _synth():
    entry.load_package()
    return importlib.import_module("nrc_b200.synthetic")


def n_mlp(enc):
    return 64 * (80 if enc == 0 else 64) + 4 * 4096 + 1024


def seeded_master(enc, n_params, visible_grid):
    """Xavier-uniform matrices (scale sqrt(6/(fan_in+fan_out)), gpu_matrix.h:284-299) from numpy PCG64(1337);
    grid U(-1e-4,1e-4) as the reference initialises it, or U(-1,1) so that the grid features are numerically visible."""
    rng = np.random.Generator(np.random.PCG64(1337 + enc))
    E = 80 if enc == 0 else 64
    parts = []
    for rows, cols in [(64, E)] + [(64, 64)] * 4 + [(16, 64)]:
        s = np.sqrt(6.0 / (rows + cols))
        parts.append(rng.uniform(-s, s, rows * cols))
    n_grid = n_params - n_mlp(enc)
    if n_grid:
        parts.append(rng.uniform(-1.0, 1.0, n_grid) * (1.0 if visible_grid else 1e-4))
    return np.concatenate(parts).astype(np.float32)


def batch(step):
    s = _synth()
    q = s.cornell_queries(16384, seed=50000 + step)
    return q, s.radiance_targets(q, seed=50000 + step)


def infer_queries():
    s = _synth()
    return np.concatenate([s.cornell_queries(N_INFER // 2, seed=77), s.uniform_queries(N_INFER // 2, seed=78)])


def load(name):
    p = os.path.join(GOLDEN_DIR, name)
    return np.load(p) if os.path.exists(p) else None
