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


# ---- NEXT rows (frame helpers): seeded inputs of tests/golden/ref_helpers.npz and ref_frame_*.npz
HELPER_W, HELPER_H = 48, 20                   # accumulate: 960 pixels
HELPER_TILES_X, HELPER_TILES_Y = 40, 25       # propagate: 1000 tiles
HELPER_RECORDS = 4096
ACC_CASES = [(0, 0), (0, 3), (2, 0), (4, 0), (5, 1), (1, 0), (3, 2)]   # (nrc::RenderMode, iterationIndex)
PERMUTE_CASES = [65536, 1000, 70000]          # numTrainingRecords (70000: clamped to 65536, nrc_helpers.cu:237)


def helper_inputs():
    """Seeded inputs for the reference's accumulate / propagate / permute kernels (nrc_helpers.cu:77-249)."""
    rng = np.random.Generator(np.random.PCG64(4711))
    n_px = HELPER_W * HELPER_H
    d = {"acc_radiance": rng.uniform(0, 2, (n_px, 3)).astype(np.float32),
         "acc_throughput": rng.uniform(0, 1, (n_px, 3)).astype(np.float32),
         "acc_output": rng.uniform(0, 1, (n_px, 4)).astype(np.float32)}
    s = _synth()
    n_tiles = HELPER_TILES_X * HELPER_TILES_Y
    ev, rec, used = s.training_chains(n_tiles, n_records=HELPER_RECORDS, seed=99, max_len=5)
    d.update(prop_ev=ev, prop_rec=rec, prop_used=used,
             prop_end_radiance=rng.uniform(0, 3, (n_tiles, 3)).astype(np.float32),
             prop_targets=rng.uniform(0, 1, (HELPER_RECORDS, 3)).astype(np.float32))
    d["perm_q"] = s.cornell_queries(65536, seed=31)
    d["perm_t"] = s.radiance_targets(d["perm_q"], seed=31)
    d["perm"] = rng.permutation(65536).astype(np.int32)      # stands for the reference's cuRAND + CUB result (NRCUtil.cu:20-35)
    return d


def digest(a):
    import hashlib
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


# One small frame through the whole NRC sequence of Device::render (Device.cpp:2495-2515): 320x240 pixels, 2x2-pixel tiles
FRAME_W, FRAME_H, FRAME_TILE = 320, 240, 2
FRAME_SEED = 5


def frame_inputs(enc):
    """Seeded frame: queries (pixels + tile tails), throughput, output image, suffix end vertices, training records,
    training queries / targets, and the permutation the reference's permute_train_data is given (this library's keyed
    bijection, so that both sides train on identical batches)."""
    s = _synth()
    entry.load_package()
    pkg = importlib.import_module("nrc_b200")
    n_px = FRAME_W * FRAME_H
    tx, ty = FRAME_W // FRAME_TILE, FRAME_H // FRAME_TILE
    n_tiles = tx * ty
    rng = np.random.Generator(np.random.PCG64(2024 + enc))
    q = np.concatenate([s.cornell_frame(FRAME_W, FRAME_H, seed=11), s.cornell_queries(n_tiles, seed=12)])
    ev, rec, used = s.training_chains(n_tiles, seed=13, max_len=3)
    tq = s.cornell_queries(65536, seed=14)
    tt = s.radiance_targets(tq, seed=14)
    tt[used:] = 0
    perm = np.array([pkg.permutation(i, FRAME_SEED) for i in range(65536)], dtype=np.int32)
    return {"queries": q, "n_px": n_px, "tiles": (tx, ty), "throughput": rng.uniform(0, 1, (n_px, 3)).astype(np.float32),
            "output": rng.uniform(0, 1, (n_px, 4)).astype(np.float32), "ev": ev, "rec": rec, "used": used,
            "train_q": tq, "train_t": tt, "perm": perm}
