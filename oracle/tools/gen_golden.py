#!/usr/bin/env python3
"""TEST TOOLING — generates the committed golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref/libjplref_sse2.so, built in place from /root/reference by oracle/Makefile) on seeded inputs.

Only data is written (inputs, outputs, hashes); no reference code. Run in the build container (needs /root/reference at
build time of oracle/_ref), commit the output:

    make -C oracle ref && python oracle/tools/gen_golden.py

Fixtures (all small):
  speaker_tris.json     triangle lists of SpeakerTriangulation::TriangulateSpeakerLayout (inc/Panning/VBAPEx.h:141-234) for the
                        14 standard 3D layouts (SpatializationTests/src/Tests/VBAP3DTest.cpp:130-154); key = mask without LFE
  lut_hashes.json       sha256 of every panner LUT (2D: 1024 x L floats; 3D: 65536 x 3 floats + 65536 x 3 bytes), shortest aperture,
                        source-layout dimensions for the source masks the reference's tests use
  primitives.npz        SinCos / acos / tan / InvSqrt / octahedral + diamond codecs / Slerp / QuatLookAt on seeded inputs
  hello_quad.npz        HelloJPLSpatial (Examples/HelloJPLSpatial/main.cpp:34-60): quad, 360 directions + the 4 printed ones
  point_gains.npz       GetSpeakerGains for seeded + special directions on quad / 5.1 / 7.1 / 5.0.2 / 7.1.4 / 9.1.6
  mdap.npz              ProcessVBAPData for (source, target) pairs incl. BASELINE config 2's mono -> 7.1.4
  crossover.npz         FourBandCrossover coefficients, impulse + noise responses (ramped gains, split)
  delay.npz             DelayLine + Thiran / linear taps
  render.npz            DirectSoundEffect::ProcessInterleaved: mono -> 7.1.4 and stereo -> 5.1, 3 blocks of 512 frames (+ a 100-frame block)
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
sys.path.insert(0, ROOT)
from oracle import refpy          # noqa: E402
from tests import util            # noqa: E402
from jplspatial_b200._masks import ChannelMask as M, EChannel as E  # noqa: E402  (pure-python constants, no native code)

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)
ref = refpy.Ref("sse2")

LAYOUTS_3D = ["Surround_5_0_2", "Surround_5_1_2", "Surround_5_0_4", "Surround_5_1_4", "Surround_7_0_2", "Surround_7_1_2", "Surround_7_0_4",
              "Surround_7_1_4", "Surround_7_0_6", "Surround_7_1_6", "Surround_9_0_4", "Surround_9_1_4", "Surround_9_0_6", "Surround_9_1_6"]
LAYOUTS_2D = ["Stereo", "LCR", "Quad", "Surround_4_1", "Surround_5_0", "Surround_5_1", "Surround_6_0", "Surround_6_1", "Surround_7_0", "Surround_7_1", "Octagonal"]
SOURCES = ["Mono", "Stereo", "LCR", "Quad", "Surround_5_0", "Surround_5_1", "Surround_7_1"]


def sha(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


# ---- speaker triangles
tris = {}
for name in LAYOUTS_3D:
    mask = getattr(M, name)
    v, t = ref.triangulate(mask)
    assert t is not None, name
    tris["0x%06x" % (mask & ~E.LFE)] = t.reshape(-1).tolist()
json.dump(tris, open(os.path.join(OUT, "speaker_tris.json"), "w"), indent=0, sort_keys=True)

# ---- LUT hashes + layout dimensions
luts = {}
for name in LAYOUTS_2D + LAYOUTS_3D:
    mask = getattr(M, name)
    p = ref.panner(mask)
    assert p.ok, name
    g, s = p.lut()
    entry = {"mask": mask, "is3d": p.is3d, "channels": p.L, "aperture_bits": int(np.float32(p.aperture).view(np.uint32)),
             "sha256": sha(g, s) if s is not None else sha(g), "layouts": {}}
    for sname in SOURCES:
        lay = p.layout(getattr(M, sname))
        if not lay.ok:
            entry["layouts"][sname] = None
            continue
        q, ids = lay.groups()
        entry["layouts"][sname] = {"channels": lay.num_channels, "vs": lay.num_virtual_sources, "rings": lay.num_rings, "per_ring": lay.samples_per_ring,
                                   "phi_bits": int(np.float32(lay.phi_term).view(np.uint32)), "offset_bits": int(np.float32(lay.offset_term).view(np.uint32)),
                                   "weight_bits": int(np.float32(lay.vs_weight).view(np.uint32)), "quats_sha256": sha(q), "ids": ids.tolist()}
    luts[name] = entry
json.dump(luts, open(os.path.join(OUT, "lut_hashes.json"), "w"), indent=1, sort_keys=True)

# ---- primitives
n = 4096
x_small = (util.uniform01(n, util.SEED + 100) * 2 - 1) * np.float32(8.0)
x_unit = util.uniform01(n, util.SEED + 101) * 2 - 1
x_unit[:8] = [-1, 1, 0, 0.5, -0.5, 0.99999994, -0.99999994, 1e-8]
x_pos = np.exp((util.uniform01(n, util.SEED + 102) * 40 - 20).astype(np.float64)).astype(np.float32)
x_tan = (util.uniform01(n, util.SEED + 103) * np.float32(1.5)).astype(np.float32)
dirs = np.concatenate([util.special_directions(), util.unit_directions(n - 26, util.SEED + 104)])
v1 = util.unit_directions(n, util.SEED + 105)
v1[:26] = dirs[:26]                      # parallel
v1[26:52] = -dirs[26:52]                 # opposite
s_, c_ = ref.sincos(x_small)
est, nr1, nr2 = ref.invsqrt(x_pos)
codes = (util.splitmix64(n, util.SEED + 106) & np.uint64(0xFFFF)).astype(np.uint32)
dia_in = dirs[:, [0, 2]].copy()
dia_in = dia_in[(np.abs(dia_in).sum(1) > 1e-3)]
p01 = util.uniform01(1024, util.SEED + 107)
np.savez_compressed(os.path.join(OUT, "primitives.npz"),
                    x_small=x_small, sin=s_, cos=c_, x_unit=x_unit, acos=ref.acos(x_unit), x_tan=x_tan, tan=ref.tan(x_tan),
                    x_pos=x_pos, rsqrt_nr1=nr1, rsqrt_nr2=nr2,
                    dirs=dirs, oct_simd=ref.oct_encode_simd(dirs), oct_scalar=ref.oct_encode_scalar(dirs),
                    codes=codes, oct_decoded=ref.oct_decode_scalar(codes),
                    dia_in=dia_in, diamond=ref.diamond_scalar(dia_in), p01=p01, from_diamond=ref.from_diamond(p01),
                    slerp_v1=v1, slerp_t=np.float32([0.0, 0.25, 0.7, 1.0]),
                    slerp_out=np.stack([ref.slerp_simd(dirs, v1, t) for t in (0.0, 0.25, 0.7, 1.0)]),
                    quat_lookat=ref.quat_lookat(dirs),
                    alpha480=np.float32(ref.smoother_alpha(480)), smoother=ref.smoother_run(10.0, 250.0, 480, 600))

# ---- Hello (config 1)
ang = np.deg2rad(np.arange(360, dtype=np.float64))
hello_dirs = np.stack([np.sin(ang), np.zeros(360), -np.cos(ang)], 1).astype(np.float32)
hello_dirs = np.concatenate([hello_dirs, np.float32([[0, 0, -1], [1, 0, 0], [0, 0, 1], [-1, 0, 0]])])
np.savez_compressed(os.path.join(OUT, "hello_quad.npz"), dirs=hello_dirs, gains=ref.panner(M.Quad).point_gains(hello_dirs))

# ---- point gains
pg = {}
pd = np.concatenate([util.special_directions(), util.unit_directions(486, util.SEED + 200)])
pg["dirs"] = pd
for name in ("Quad", "Surround_5_1", "Surround_7_1", "Stereo", "Surround_5_0_2", "Surround_7_1_4", "Surround_9_1_6"):
    pg[name] = ref.panner(getattr(M, name)).point_gains(pd)
np.savez_compressed(os.path.join(OUT, "point_gains.npz"), **pg)

# ---- MDAP
md = {}
params = util.pan_params(512, util.SEED + 300)
md["params"] = params
for sname, tname in (("Mono", "Surround_7_1_4"), ("Stereo", "Surround_7_1_4"), ("Surround_5_1", "Surround_9_1_6"), ("Mono", "Surround_9_1_6"), ("Quad", "Surround_5_0_2"),
                     ("Mono", "Quad"), ("Stereo", "Surround_7_1"), ("Mono", "Stereo"), ("Surround_5_0", "Surround_5_1"), ("LCR", "Surround_7_1_4")):
    p = ref.panner(getattr(M, tname))
    lay = p.layout(getattr(M, sname))
    assert lay.ok, (sname, tname)
    md[f"{sname}__{tname}"] = p.mdap(lay, params)
vs = np.concatenate([util.unit_directions(64, util.SEED + 301), util.uniform01(64, util.SEED + 302)[:, None]], 1).astype(np.float32)
md["vs"] = vs
md["vs_Surround_7_1_4"] = ref.panner(M.Surround_7_1_4).process_virtual_sources(vs)
md["vs_Quad"] = ref.panner(M.Quad).process_virtual_sources(vs)
np.savez_compressed(os.path.join(OUT, "mdap.npz"), **md)

# ---- crossover
xo = ref.crossover(48000.0)
imp = np.zeros(1024, np.float32); imp[0] = 1.0
noise = util.white_noise((1024,), util.SEED + 400)
g0 = np.float32([0.9, 0.5, 0.25, 0.1]); g1 = np.float32([0.2, 1.0, 0.7, 0.6])
cz = {"coeffs": xo.coeffs(), "impulse": imp, "noise": noise, "g0": g0, "g1": g1}
xo.reset(); cz["impulse_split"] = xo.split(imp)
xo.reset(); cz["noise_ramp_a"] = xo.process_ramp(noise[:480], g0, g1); cz["noise_ramp_b"] = xo.process_ramp(noise[480:512], g1, g1); cz["noise_ramp_c"] = xo.process_ramp(noise[512:], g1, g0)
np.savez_compressed(os.path.join(OUT, "crossover.npz"), **cz)

# ---- delay line
dl = ref.delay(1000)
sig = util.white_noise((3000,), util.SEED + 500)
dl.push(sig[:2000])
dz = {"signal": sig, "size": np.uint32(dl.size())}
dz["int_delays"] = np.uint32([0, 1, 7, 100, 999, 1023])
dz["int_reads"] = np.float32([dl.read_int(int(d)) for d in dz["int_delays"]])
fd = (util.uniform01(256, util.SEED + 501) * 900).astype(np.float32); fd[:4] = [0.0, 1.0, 100.0, 0.25]
dz["frac_delays"] = fd
dz["thiran"] = np.float32([dl.tap_thiran(float(d)) for d in fd])
dz["linear"] = np.float32([dl.tap_linear(float(d)) for d in fd])
np.savez_compressed(os.path.join(OUT, "delay.npz"), **dz)

# ---- DirectSoundEffect
rz = {}
for tag, smask, tmask, nsrc in (("mono_714", M.Mono, M.Surround_7_1_4, 3), ("stereo_51", M.Stereo, M.Surround_5_1, 2)):
    p = ref.panner(tmask); lay = p.layout(smask)
    C, L = lay.num_channels, p.L
    blocks = [512, 512, 512, 100]
    prm = util.pan_params(nsrc * (len(blocks) + 1), util.SEED + 600, edge=False)
    mix = p.mdap(lay, prm).reshape(len(blocks) + 1, nsrc, C * L)
    band = (0.1 + 0.9 * util.uniform01((len(blocks) + 1) * nsrc * 4, util.SEED + 601)).reshape(len(blocks) + 1, nsrc, 4).astype(np.float32)
    delay = ((1.0 + 60.0 * util.uniform01((len(blocks) + 1) * nsrc, util.SEED + 602)) / 343.2).reshape(len(blocks) + 1, nsrc).astype(np.float32)
    delay[:, 0] = np.float32(0.004)                      # a static source
    audio = [util.white_noise((nsrc, f, C), util.SEED + 610 + i) for i, f in enumerate(blocks)]
    fx = [ref.dse(48000.0, C, L, band[0, i], float(delay[0, i]), mix[0, i]) for i in range(nsrc)]
    rz[f"{tag}_mix"] = mix; rz[f"{tag}_band"] = band; rz[f"{tag}_delay"] = delay
    for b, f in enumerate(blocks):
        out = np.zeros((nsrc, f, L), np.float32)
        for i in range(nsrc):
            fx[i].update(band[b + 1, i], float(delay[b + 1, i]), mix[b + 1, i])
            fx[i].process(audio[b][i], out[i], f)
        rz[f"{tag}_in{b}"] = audio[b]; rz[f"{tag}_out{b}"] = out
np.savez_compressed(os.path.join(OUT, "render.npz"), **rz)

# ---- ReverbBus + ERBus (BASELINE config 5)
vz = {}
FDN_DELAYS = np.array([2927, 2593, 2273, 3697, 1877, 3877, 2477, 3461, 1609, 3779, 3541, 4259, 1669, 3539, 3637, 4013], np.float32)
vz["db"] = (-60.0 * util.uniform01(64, util.SEED + 700)).astype(np.float32)
vz["db_gain"] = ref.db_to_gain(vz["db"]).reshape(-1)
rv = ref.reverb(48000.0)
rt = np.array([[0.4, 1.5, 2.5, 0.8], [0.01, 20.0, 1.0, 0.2]], np.float32)
vz["rev_rt60"] = rt
vz["rev_in"] = (util.white_noise((7, 512), util.SEED + 701) * np.float32(0.5)); vz["rev_in"][0, 0] += 1.0; vz["rev_in"][5:] = 0.0
outs = []
for b in range(7):
    if b == 0: rv.set_rt60(rt[0])
    if b == 4: rv.set_rt60(rt[1])
    o = np.zeros((512, 4), np.float32); rv.process(vz["rev_in"][b].reshape(512, 1), o, 512); outs.append(o)
vz["rev_out"] = np.stack(outs)
eb = ref.erbus(48000.0, 4)
sched = [[1, 2, 3], [1, 2, 3], [2, 3, 7], [2, 3, 7], [3, 7, 1], [3, 7, 1]]
vz["er_in"] = util.white_noise((6, 512, 2), util.SEED + 702)
eouts, ntaps = [], []
for b, ids in enumerate(sched):
    u = util.uniform01(len(ids) * 9, util.SEED + 710 + b).reshape(len(ids), 9)
    fg = (0.1 + 0.9 * u[:, :4]).astype(np.float32); d = (0.002 + 0.05 * np.array(ids, np.float32) / 10.0 + 0.0002 * u[:, 4]).astype(np.float32); g = u[:, 5:9].astype(np.float32)
    vz[f"er_fg{b}"] = fg; vz[f"er_delay{b}"] = d; vz[f"er_ids{b}"] = np.array(ids, np.uint32); vz[f"er_gains{b}"] = g
    eb.set_taps(fg, d, ids, g); ntaps.append(eb.num_taps())
    o = np.zeros((512, 4), np.float32); eb.process(vz["er_in"][b], o, 512, n_in=2); eouts.append(o)
vz["er_out"] = np.stack(eouts); vz["er_num_taps"] = np.array(ntaps, np.uint32)
np.savez_compressed(os.path.join(OUT, "reverb_er.npz"), **vz)

total = sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT))
print("wrote", sorted(os.listdir(OUT)), f"{total / 1024:.0f} KiB")
