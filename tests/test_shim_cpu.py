"""The drop-in boundary as source code: include/jplg.h is a plain-C header, and include/jplg_shim.hpp — the adapter
INTEGRATION.md hands to a JPLSpatial maintainer — compiles against the reference's REAL headers (JPL::StandardPanner3D::PanUpdateData,
JPL::ChannelMap, JPL::simd, DirectSoundEffect's signatures) and links against libjplg.so. The compiled program (oracle/_ref/shim_check, built
from oracle/ref/shim_check.cpp by oracle/Makefile) runs here on a host-only context; tests/test_gpu_shim.py runs the same binary on the device."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")
REF_INC = "/root/reference/Spatialization/include"
CHECK = os.path.join(ROOT, "oracle", "_ref", "shim_check")


def test_jplg_h_is_a_plain_c_header(tmp_path):
    """cgo / JNI / ctypes bind a C header: jplg.h alone, as C11, with every warning an error"""
    tu = tmp_path / "only_c.c"
    tu.write_text('#include "jplg.h"\n'
                  "int main(void) { jplg_renderer_config cfg; jplg_context c = 0; (void)c; cfg.num_sources = 0; return (int)cfg.num_sources + (int)sizeof(jplg_frontend_io) * 0; }\n")
    for cc, std in (("gcc", "-std=c11"), ("gcc", "-std=c99"), ("g++", "-std=c++17")):
        if shutil.which(cc) is None:
            pytest.skip(f"{cc} not installed")
        args = [cc, std, "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", INC] + (["-x", "c++"] if cc == "g++" else []) + [str(tu)]
        r = subprocess.run(args, capture_output=True, text=True)
        assert r.returncode == 0, f"{' '.join(args)}\n{r.stderr}"


@pytest.mark.skipif(not os.path.isdir(REF_INC), reason="the reference tree is not on this machine")
def test_shim_header_compiles_against_the_reference_headers(tmp_path):
    tu = tmp_path / "engine_side.cpp"
    tu.write_text('''
#include <JPLSpatial/Panning/VBAPanning2D.h>
#include <JPLSpatial/Panning/VBAPanning3D.h>
#include "jplg_shim.hpp"
// an engine's call sites, 2D and 3D panner types alike
template<class P> jplg_status Tick(JPLG::BatchedPanner<P>& pan, JPLG::DirectSoundBatch& batch, std::span<const typename P::PanUpdateData> dirty,
                                   std::span<float> maps, std::span<const float> gains, std::span<const float> delays)
{
    if (jplg_status st = pan.ProcessVBAPData(dirty, maps)) return st;
    return batch.UpdateParameters(gains, delays, maps);
}
template jplg_status Tick<JPL::StandardPanner2D>(JPLG::BatchedPanner<JPL::StandardPanner2D>&, JPLG::DirectSoundBatch&, std::span<const JPL::StandardPanner2D::PanUpdateData>, std::span<float>, std::span<const float>, std::span<const float>);
template jplg_status Tick<JPL::StandardPanner3D>(JPLG::BatchedPanner<JPL::StandardPanner3D>&, JPLG::DirectSoundBatch&, std::span<const JPL::StandardPanner3D::PanUpdateData>, std::span<float>, std::span<const float>, std::span<const float>);
void Callback(JPLG::DirectSoundBatch& batch, std::span<const float> in, std::span<float> bus, JPL::uint32 numFrames) { batch.ProcessInterleaved(in, bus, numFrames); }
''')
    args = ["g++", "-std=c++20", "-fsyntax-only", "-Wall", "-Wextra", "-Wno-unused-parameter", "-Wno-unused-variable", "-DJPL_SIMD_CONFIGURED=1", "-DNDEBUG",
            "-I", REF_INC, "-I", INC, str(tu)]
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0, f"{' '.join(args)}\n{r.stderr[-4000:]}"


def test_shim_program_runs_on_a_host_only_context():
    """the linked program: reference objects and shim objects side by side; without a device everything that needs none works and every compute
    call fails with JPLG_ERR_NO_DEVICE"""
    if not os.path.exists(CHECK):
        pytest.skip("oracle/_ref/shim_check not built (needs /root/reference at build time)")
    r = subprocess.run([CHECK, "-1"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failure(s) (host-only)" in r.stdout and "FAIL" not in r.stdout
