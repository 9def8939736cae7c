"""Seeded synthetic RadianceQuery / target buffers that stand in for the OptiX path tracer's output.

Shape of the data follows the reference's Cornell scene: hit points on the six walls of a box
of half-extent 10 scaled by 0.005 (nrc/shaders/hit.cu:597), directions/normals as raw spherical angles
(theta in [0, pi], phi in (-pi, pi]; nrc/shaders/shader_common.h:320-333), roughness in [0,1]^2, diffuse albedo
by wall, specular mostly 0.  Record layout = nrc::RadianceQuery (neural_radiance_caching.h:100-118):
[pos(3) | dir(theta,phi) | normal(theta,phi) | roughness(2) | diffuse(3) | specular(3)]  = 15 float32.
"""
import numpy as np


def _sph(v):
    """cartesianToSphericalUnitVector: theta = 2*asin(|v - z|/2), phi = atan2(y, x)."""
    d = v - np.array([0.0, 0.0, 1.0])
    theta = 2.0 * np.arcsin(np.clip(0.5 * np.linalg.norm(d, axis=-1), 0.0, 1.0))
    phi = np.arctan2(v[..., 1], v[..., 0])
    return theta, phi


def cornell_queries(n, seed=1337):
    """n x 15 float32 queries (C-contiguous)."""
    rng = np.random.Generator(np.random.Philox(seed))
    wall = rng.integers(0, 6, n)
    axis, sign = wall // 2, (wall % 2) * 2.0 - 1.0
    uv = rng.uniform(-10.0, 10.0, (n, 2))
    pos = np.empty((n, 3))
    nrm = np.zeros((n, 3))
    for a in range(3):
        m = axis == a
        o = [i for i in range(3) if i != a]
        pos[m, a] = sign[m] * 10.0
        pos[np.ix_(m, o)] = uv[m]
        nrm[m, a] = -sign[m]
    # cosine-weighted hemisphere direction about the normal
    u1, u2 = rng.random(n), rng.random(n)
    r, ph = np.sqrt(u1), 2.0 * np.pi * u2
    lx, ly, lz = r * np.cos(ph), r * np.sin(ph), np.sqrt(1.0 - u1)
    t = np.where(np.abs(nrm[:, [0]]) > 0.5, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
    b1 = np.cross(nrm, t)
    b1 /= np.linalg.norm(b1, axis=1, keepdims=True)
    b2 = np.cross(nrm, b1)
    dirs = lx[:, None] * b1 + ly[:, None] * b2 + lz[:, None] * nrm
    q = np.empty((n, 15), dtype=np.float32)
    q[:, 0:3] = pos * 0.005
    q[:, 3], q[:, 4] = _sph(dirs)
    q[:, 5], q[:, 6] = _sph(nrm)
    q[:, 7:9] = rng.random((n, 2))
    diffuse = np.full((n, 3), 0.8)
    diffuse[wall == 0] = (0.8, 0.1, 0.1)   # red wall
    diffuse[wall == 1] = (0.1, 0.8, 0.1)   # green wall
    q[:, 9:12] = diffuse
    spec = np.zeros((n, 3))
    m = rng.random(n) < 0.1
    spec[m] = rng.random((int(m.sum()), 3))
    q[:, 12:15] = spec
    return np.ascontiguousarray(q)


def cornell_frame(width, height, seed=1337, n_tail=0):
    """width*height x 15 float32 queries in PIXEL ORDER (row-major), the way Device::render leaves them in
    radianceQueriesInference: one primary hit per pixel of a pinhole camera looking into the Cornell box through its
    open face (camera on the +z axis, 1 jittered sample per pixel), followed by n_tail incoherent queries (the path
    tails of the training tiles).  Neighbouring pixels hit neighbouring surface points, which is what makes the hash-grid
    gathers of a real frame cache-friendly; cornell_queries() is the incoherent worst case."""
    rng = np.random.Generator(np.random.Philox(seed + 3))
    n = width * height
    jx, jy = rng.random(n, dtype=np.float32), rng.random(n, dtype=np.float32)
    px = (np.arange(n, dtype=np.int64) % width).astype(np.float32)
    py = (np.arange(n, dtype=np.int64) // width).astype(np.float32)
    u = ((px + jx) / width * 2.0 - 1.0).astype(np.float64) * 0.5
    v = ((py + jy) / height * 2.0 - 1.0).astype(np.float64) * (0.5 * height / width)
    d = np.stack([u, v, -np.ones(n)], axis=1)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = np.array([0.0, 0.0, 30.0])
    # nearest hit among the five walls x = +-10, y = +-10, z = -10 (the face z = +10 is open)
    tbest = np.full(n, np.inf)
    wall = np.zeros(n, dtype=np.int64)
    for w, (a, sgn) in enumerate([(0, -1.0), (0, 1.0), (1, -1.0), (1, 1.0), (2, -1.0)]):
        with np.errstate(divide="ignore", invalid="ignore"):
            t = (sgn * 10.0 - o[a]) / d[:, a]
        hit = o[None, :] + t[:, None] * d
        ok = (t > 0) & np.all(np.abs(hit) <= 10.0 + 1e-9, axis=1) & (t < tbest)
        tbest = np.where(ok, t, tbest)
        wall = np.where(ok, w, wall)
    pos = o[None, :] + tbest[:, None] * d
    nrm = np.zeros((n, 3))
    axis, sgn = np.array([0, 0, 1, 1, 2])[wall], np.array([-1.0, 1.0, -1.0, 1.0, -1.0])[wall]
    nrm[np.arange(n), axis] = -sgn
    q = np.empty((n, 15), dtype=np.float32)
    q[:, 0:3] = pos * 0.005
    q[:, 3], q[:, 4] = _sph(-d)
    q[:, 5], q[:, 6] = _sph(nrm)
    q[:, 7] = (0.3 + 0.05 * wall).astype(np.float32)            # per-material roughness
    q[:, 8] = 0.5
    diffuse = np.full((n, 3), 0.8)
    diffuse[wall == 0] = (0.8, 0.1, 0.1)   # red wall
    diffuse[wall == 1] = (0.1, 0.8, 0.1)   # green wall
    q[:, 9:12] = diffuse
    q[:, 12:15] = 0.0
    if n_tail:
        q = np.concatenate([q, cornell_queries(n_tail, seed=seed + 5)], axis=0)
    return np.ascontiguousarray(q)


def uniform_queries(n, seed=1337):
    """All 15 inputs ~ U[0,1): the one-blob encoding in its non-degenerate regime."""
    rng = np.random.Generator(np.random.Philox(seed + 7))
    return np.ascontiguousarray(rng.random((n, 15), dtype=np.float32))


def radiance_targets(q, seed=1337):
    """Smooth analytic radiance + 1 % multiplicative noise, n x 3 float32, >= 0."""
    rng = np.random.Generator(np.random.Philox(seed + 13))
    q = q.astype(np.float64)
    base = (0.5 + 0.5 * np.cos(40.0 * q[:, 0]) * np.cos(40.0 * q[:, 1])) * (1.0 + q[:, 5] / np.pi)
    t = q[:, 9:12] * base[:, None] * (1.0 + 0.01 * rng.standard_normal((q.shape[0], 1)))
    return np.ascontiguousarray(np.maximum(t, 0.0).astype(np.float32))


def training_chains(n_tiles, n_records=65536, seed=1337, max_len=2):
    """Synthetic stand-in for what the path tracer leaves for radiance propagation (nrc/shaders/hit.cu): one
    nrc::TrainingSuffixEndVertex per tile pointing at a chain of 0..max_len nrc::TrainingRecord's linked by propTo.
    Returns (end_vertices int32[n_tiles, 4], records int32[n_records, 7], records_used); float fields are stored as
    their bit patterns (layouts: include/nrc_b200.h)."""
    rng = np.random.Generator(np.random.Philox(seed + 29))
    length = rng.integers(0, max_len + 1, n_tiles)
    start = np.concatenate([[0], np.cumsum(length)[:-1]])
    length = np.where(start + length <= n_records, length, 0)
    used = int((start + length).max()) if n_tiles else 0
    rec = np.zeros((n_records, 7), dtype=np.int32)
    rec[:, 0] = -1
    ids = np.arange(n_records)
    owner = np.searchsorted(start + length, ids, side="right")          # tile owning record i (if i < used)
    valid = (ids < used) & (owner < n_tiles)
    last = np.zeros(n_records, dtype=bool)
    last[valid] = ids[valid] == (start + length)[owner[valid]] - 1
    rec[valid & ~last, 0] = ids[valid & ~last] + 1
    rec[:, 1:4] = rng.uniform(0.0, 1.1, (n_records, 3)).astype(np.float32).view(np.int32)
    rec[valid, 5] = owner[valid]
    ev = np.zeros((n_tiles, 4), dtype=np.int32)
    ev[:, 0] = np.where(length > 0, start, -1)
    ev[:, 1] = (rng.random(n_tiles) > 1.0 / 16.0).astype(np.float32).view(np.int32)   # TRAIN_UNBIASED_RATIO = 1/16
    ev[:, 3] = np.arange(n_tiles)
    return ev, rec, used
