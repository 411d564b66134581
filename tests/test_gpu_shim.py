"""The reference-side adapter (include/jplg_shim.hpp) on the device: oracle/_ref/shim_check holds the unmodified reference's panner and
DirectSoundEffect next to JPLG::BatchedPanner / JPLG::DirectSoundBatch / JPLG::Mixing, feeds both the same reference-typed arguments
(JPL::StandardPanner3D::PanUpdateData, PanUpdateDataWithOrientation, JPL::simd gains) and compares bit for bit. Built here by oracle/Makefile;
the binary travels to the GPU box like the other prebuilt checker files."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECK = os.path.join(ROOT, "oracle", "_ref", "shim_check")


def test_shim_matches_reference_objects_bit_for_bit(gpu_ctx):
    if not os.path.exists(CHECK):
        pytest.skip("oracle/_ref/shim_check not built (needs /root/reference at build time)")
    r = subprocess.run([CHECK, str(gpu_ctx.device)], capture_output=True, text=True, timeout=300)
    print(r.stdout)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "shim_check: 0 failure(s)" in r.stdout and "FAIL" not in r.stdout
    for line in ("ProcessVBAPData(PanUpdateData): shim == reference", "ProcessVBAPData(PanUpdateDataWithOrientation)", "one-source batches == reference effects",
                 "Mixing::Add == JPL::Add", "Mixing::DownmixToMono == JPL::DownmixToMono"):
        assert line in r.stdout, line
