"""Synthetic point clouds for tests, the oracle harness and bench.py (input generation only).

TEST INFRASTRUCTURE + input generation only (numpy, no reference code, no compute path).
Shapes follow SURVEY.md §8(d): points sampled on analytic surfaces so that kNN patches look like
surface patches (unit sphere, torus R=1 r=0.35, unit-cube surface), fp32, seeded.
"""
import numpy as np

SHAPES = ("sphere", "torus", "cube")


def make_cloud(n: int, seed: int = 2021, shape: str = "sphere") -> np.ndarray:
    """Return one [n, 3] float32 cloud on the named surface."""
    rng = np.random.default_rng(seed)
    if shape == "sphere":
        v = rng.standard_normal((n, 3))
        v /= np.linalg.norm(v, axis=1, keepdims=True)
    elif shape == "torus":
        u = rng.uniform(0.0, 2.0 * np.pi, n)
        w = rng.uniform(0.0, 2.0 * np.pi, n)
        R, r = 1.0, 0.35
        v = np.stack([(R + r * np.cos(w)) * np.cos(u), (R + r * np.cos(w)) * np.sin(u), r * np.sin(w)], 1)
    elif shape == "cube":
        face = rng.integers(0, 6, n)
        ab = rng.uniform(-1.0, 1.0, (n, 2))
        v = np.empty((n, 3))
        axis = face // 2
        sign = np.where(face % 2 == 0, 1.0, -1.0)
        for a in range(3):
            m = axis == a
            o = [i for i in range(3) if i != a]
            v[m, a] = sign[m]
            v[m, o[0]] = ab[m, 0]
            v[m, o[1]] = ab[m, 1]
    elif shape == "uniform":
        v = rng.uniform(-1.0, 1.0, (n, 3))
    elif shape == "lattice":
        # integer lattice: exact distance ties everywhere (kNN / Chamfer tie stress)
        side = int(np.ceil(n ** (1.0 / 3.0)))
        g = np.stack(np.meshgrid(*[np.arange(side)] * 3, indexing="ij"), -1).reshape(-1, 3).astype(np.float64)
        v = g[rng.permutation(len(g))[:n]]
    else:
        raise ValueError(shape)
    return np.ascontiguousarray(v, dtype=np.float32)


def make_clouds(b: int, n: int, seed: int = 2021) -> np.ndarray:
    """[b, n, 3]: mixed shapes, seeds seed, seed+1, ... (config 2 of BASELINE.json)."""
    return np.stack([make_cloud(n, seed + i, SHAPES[i % 3]) for i in range(b)], 0)


def normalize_patches(p: np.ndarray) -> np.ndarray:
    """Per-item centroid / max-radius normalisation in fp32 (same arithmetic order as the reference's
    normalize_pc, /root/reference/modules/utils/patch.py:168-178)."""
    p = p.astype(np.float32)
    c = p.mean(axis=1, keepdims=True, dtype=np.float32)
    q = p - c
    r = np.sqrt((q * q).sum(-1, keepdims=True, dtype=np.float32)).max(axis=1, keepdims=True)
    return (q / r).astype(np.float32)


def make_patches(bp: int, n: int = 256, seed: int = 2021, shape: str = "uniform") -> np.ndarray:
    """[bp, n, 3] per-patch normalised patches (config 4 microbench style)."""
    rng = np.random.default_rng(seed)
    if shape == "uniform":
        p = rng.uniform(-1.0, 1.0, (bp, n, 3)).astype(np.float32)
    elif shape == "surface":
        # a curved surface patch: z = quadratic(x, y) with random curvature + small noise
        xy = rng.uniform(-1.0, 1.0, (bp, n, 2))
        k = rng.uniform(-0.6, 0.6, (bp, 1, 3))
        z = k[..., 0] * xy[..., 0] ** 2 + k[..., 1] * xy[..., 1] ** 2 + k[..., 2] * xy[..., 0] * xy[..., 1]
        p = np.concatenate([xy, z[..., None]], -1).astype(np.float32)
    else:
        raise ValueError(shape)
    return normalize_patches(p)
