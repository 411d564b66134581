#!/usr/bin/env python3
"""TEST TOOLING — golden fixture for the scene front-end, written by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref/libjplref_sse2.so) on seeded inputs. Only data is written. Companion of gen_golden.py (kept separate so the older
fixtures are not rewritten):

    make -C oracle ref && python oracle/tools/gen_golden_frontend.py        -> tests/golden/frontend.npz

Contents: DirectPathService::ProcessDirectPath, ProcessAngleAttenuationImpl, AttenuationBaseModel / AttenuationCurve::Evaluate,
AirAbsorption::CacheParameters (default + two environments) and ComputeForDistance + dBToGain, SpeedOfSound, SanitizeSourceDir,
GetSpreadFromSourceSize (both flavours), ProcessVBAPData with orientation (mono -> 7.1.4)."""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
sys.path.insert(0, ROOT)
from oracle import refpy          # noqa: E402
from tests import util            # noqa: E402
from tests.frontend_inputs import make_positions, CURVE_A, CURVE_B, MODELS, ENVS  # noqa: E402
from jplspatial_b200._masks import ChannelMask as M  # noqa: E402

ref = refpy.Ref("sse2")
z = {}
n = 512
src = make_positions(n, util.SEED + 100); lis = make_positions(3, util.SEED + 200, 10.0)
idx = (util.splitmix64(n, util.SEED + 300) % np.uint64(3)).astype(np.uint32)
src[0, :3] = lis[idx[0], :3]
src[1, :3] = lis[idx[1], :3] + np.float32(5.0) * lis[idx[1], 3:6]
src[2, :3] = lis[idx[2], :3] - np.float32(5.0) * lis[idx[2], 3:6]
z["src"], z["lis"], z["idx"] = src, lis, idx
z["direct"] = ref.direct_path(src, lis, idx)
z["dots"] = util.white_noise((n,), util.SEED + 400)
z["cones"] = np.sort((util.uniform01(2 * n, util.SEED + 401) * np.float32(2 * math.pi)).reshape(n, 2), axis=1).astype(np.float32)
z["cone_factor"] = ref.angle_attenuation(0, z["dots"], z["cones"])
z["dist"] = (util.uniform01(n, util.SEED + 500) * np.float32(120.0) - np.float32(5.0)).astype(np.float32)
for k, m in enumerate(MODELS):
    z[f"model{k}"] = ref.attenuation_model(*m, z["dist"])
z["curve_a"] = ref.attenuation_curve(np.array(CURVE_A, np.float32), z["dist"])
z["curve_b"] = ref.attenuation_curve(np.array(CURVE_B, np.float32), z["dist"])
z["bands"] = ref.default_bands()
z["far"] = (util.uniform01(n, util.SEED + 600) ** 2 * np.float32(3000.0)).astype(np.float32)
for k, env in enumerate(ENVS):
    c = ref.air_cache(None, *env)
    z[f"air_cache{k}"] = np.concatenate([c[key] for key in ("band_centers", "lower_third_centers", "center_db", "lower_third_db", "weighted_db",
                                                              "decay_slope", "distance_to_knee", "linearized_loss_factor")])
    z[f"air_knee{k}"] = c["has_knee"]
    z[f"air_db{k}"], z[f"air_gain{k}"] = ref.air_for_distance(z["far"], None, *env)
    z[f"sos{k}"] = np.float32(ref.speed_of_sound(env[0]))
z["inv_sos"] = np.float32(ref.inv_speed_of_sound())
dirs = z["direct"][:, 3:6].copy(); dirs[5] = [0, 1, 0]; dirs[6] = [0, -1, 0]; dirs[7] = [1e-7, -0.99999994, 0]
z["san_in"], z["san_out"] = dirs, ref.sanitize_source_dir(dirs)
z["sizes"] = (util.uniform01(n, util.SEED + 703) * np.float32(20.0)).astype(np.float32)
for f in (0, 1):
    z[f"spread{f}"] = ref.spread_from_size(f, z["sizes"], z["direct"][:, 0])
op = np.zeros((256, 11), np.float32)
op[:, :3] = z["direct"][:256, 3:6]; op[:, 3] = util.uniform01(256, util.SEED + 903); op[:, 4] = util.uniform01(256, util.SEED + 904); op[:, 5:11] = z["direct"][:256, 6:12]
rp = ref.panner(M.Surround_7_1_4)
z["oriented_params"], z["oriented_mix"] = op, rp.mdap_oriented(rp.layout(M.Mono), op)
path = os.path.join(ROOT, "tests", "golden", "frontend.npz")
np.savez_compressed(path, **z)
print("wrote", path, os.path.getsize(path), "bytes")
