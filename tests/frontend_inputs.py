"""Shared seeded inputs of the front-end tests and of oracle/tools/gen_golden_frontend.py."""
import numpy as np

from tests import util

# Curve::EType (inc/Math/Curves.h:40-52): 0 Constant, 1 Linear, 2 SineFadeIn, 3 SineFadeOut, 4 SCurve, 5 Log1.41, 6 Log3, 7 Exp1.41, 8 Exp3
CURVE_A = [(0.0, 1.0, 1), (10.0, 0.5, 4), (20.0, 0.25, 0), (40.0, 0.1, 1), (100.0, 0.0, 1)]
CURVE_B = [(50.0, 0.2, 8), (0.0, 1.0, 2), (5.0, 0.9, 3), (12.0, 0.7, 5), (25.0, 0.5, 6), (35.0, 0.3, 7), (80.0, 0.05, 1)]   # unsorted on purpose
# AttenuationModel, MinDistance, MaxDistance, Rolloff
MODELS = [(0, 1.0, 50.0, 1.0), (1, 1.0, 50.0, 1.0), (1, 2.0, 80.0, 0.5), (2, 1.0, 100.0, 1.0), (3, 1.0, 100.0, 2.0), (1, 10.0, 5.0, 1.0)]
# temperature C, relative humidity %, pressure kPa
ENVS = [(20.0, 50.0, 101.325), (5.0, 80.0, 95.0), (35.0, 15.0, 101.325)]


def make_positions(n, seed, spread_m=60.0):
    """[n,9] location, up, forward: random locations, random orthonormal up / forward."""
    loc = (util.white_noise((n, 3), seed) * np.float32(spread_m)).astype(np.float32)
    fwd = util.unit_directions(n, seed + 1)
    tmp = util.unit_directions(n, seed + 2)
    up = np.cross(fwd.astype(np.float64), tmp.astype(np.float64))
    up /= np.linalg.norm(up, axis=1, keepdims=True)
    return np.ascontiguousarray(np.concatenate([loc, up.astype(np.float32), fwd], axis=1))
