"""Deterministic synthetic inputs (splitmix64) shared by tests, smoke() and bench.py."""
import numpy as np

SEED = 0x9E3779B97F4A7C15
_M = (1 << 64) - 1


def splitmix64(n, seed=SEED):
    """n uint64 values of the splitmix64 stream started at `seed` (vectorised)."""
    with np.errstate(over="ignore"):
        idx = np.arange(1, n + 1, dtype=np.uint64)
        z = np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform01(n, seed=SEED):
    """float32 in [0,1) with 24 random bits (exact in float32)."""
    return ((splitmix64(n, seed) >> np.uint64(40)).astype(np.float64) / float(1 << 24)).astype(np.float32)


def unit_directions(n, seed=SEED):
    """Rejection-sample the unit ball (0.01 < |v|^2 <= 1), normalise with exact float32 1/sqrt."""
    out = np.zeros((0, 3), np.float32)
    k = 0
    while out.shape[0] < n:
        u = uniform01(3 * 2 * n, seed + 7919 * k).reshape(-1, 3) * np.float32(2.0) - np.float32(1.0)
        l2 = (u[:, 0] * u[:, 0] + u[:, 1] * u[:, 1] + u[:, 2] * u[:, 2]).astype(np.float32)
        u = u[(l2 > 0.01) & (l2 <= 1.0)]
        l2 = (u[:, 0] * u[:, 0] + u[:, 1] * u[:, 1] + u[:, 2] * u[:, 2]).astype(np.float32)
        inv = (np.float32(1.0) / np.sqrt(l2, dtype=np.float32)).astype(np.float32)
        out = np.concatenate([out, (u * inv[:, None]).astype(np.float32)])
        k += 1
    return np.ascontiguousarray(out[:n])


def special_directions():
    """6 axes + 26 cube directions (SpatializationTests/src/Tests/VBAP3DTest.cpp:453-486), normalised."""
    d = []
    for x in (-1, 0, 1):
        for y in (-1, 0, 1):
            for z in (-1, 0, 1):
                if x or y or z:
                    v = np.array([x, y, z], np.float32)
                    l2 = np.float32(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])
                    d.append(v * (np.float32(1.0) / np.sqrt(l2, dtype=np.float32)))
    return np.ascontiguousarray(np.array(d, np.float32))


def pan_params(n, seed=SEED, edge=True):
    """[n,5] = {dir, focus, spread}; the first rows carry forced edge values {0, 0.5, 1} and special directions."""
    p = np.zeros((n, 5), np.float32)
    p[:, :3] = unit_directions(n, seed)
    p[:, 3] = uniform01(n, seed + 1)
    p[:, 4] = uniform01(n, seed + 2)
    if edge:
        sd = special_directions()
        vals = [0.0, 0.5, 1.0]
        k = 0
        for d in sd:
            for f in vals:
                for s in vals:
                    if k < n:
                        p[k, :3] = d
                        p[k, 3] = f
                        p[k, 4] = s
                        k += 1
    return p


def white_noise(shape, seed=SEED):
    n = int(np.prod(shape))
    return (uniform01(n, seed) * np.float32(2.0) - np.float32(1.0)).reshape(shape)


def bits_equal(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def count_bit_mismatch(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return int(np.count_nonzero(a.view(np.uint32) != b.view(np.uint32)))
