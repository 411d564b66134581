"""jplg_reverb_bank_process_no_alloc (include/jplg.h): the allocation-free variant of jplg_reverb_bank_process — no counterpart in the reference.
Its mono down-mix scratch is a file-scope device array (jplg_reverb_kernels.cu) instead of the context's cudaMalloc'ed
scratch; the kernels and their order are the same, so the outputs must be bit-equal to the existing entry point's (the bar of tests/test_gpu_reverb.py).

Collected last on purpose; passes alone in a fresh process and after the other tests: it only needs its own two banks."""
import numpy as np
import pytest

import jplspatial_b200 as J
from tests import util

pytestmark = pytest.mark.gpu
FS = 48000.0


def _banks(ctx, n, L):
    rt60 = (0.3 + 2.7 * util.uniform01(n * 4, 9100 + n).reshape(n, 4)).astype(np.float32)
    a, b = J.ReverbBusBatch(ctx, n, FS, L), J.ReverbBusBatch(ctx, n, FS, L)
    a.SetRT60(rt60); b.SetRT60(rt60)
    return a, b


@pytest.mark.parametrize("n_in,frames", [(2, 480), (6, 256)])
def test_no_alloc_equals_process_on_consecutive_calls(gpu_ctx, n_in, frames):
    """two banks with the same parameters, one driven through each entry point, compared after every one of several calls in a row (the second and later
    calls find the previous call's down-mix in the shared scratch and the FDN state of the call before)"""
    import torch
    n, L, blocks = 48, 12, 12
    dev = torch.device("cuda", 0)
    stream = gpu_ctx.torch_stream()
    x = (util.white_noise((blocks, n, frames, n_in), 4242 + n_in) * np.float32(0.5)).astype(np.float32)
    x[blocks // 2:, n // 2:] = 0.0          # some instances ring out: the scratch then holds zeros where the call before left samples
    old, new = _banks(gpu_ctx, n, L)
    with torch.cuda.stream(stream):
        dx = torch.from_numpy(x).to(dev)
        want = torch.zeros((blocks, n, frames, L), dtype=torch.float32, device=dev)
        got = torch.zeros_like(want)
        gpu_ctx.synchronize()
        for b in range(blocks):
            old.ProcessInterleaved(dx[b], want[b], frames, in_channels=n_in)
            new.ProcessInterleavedNoAlloc(dx[b], got[b], frames, in_channels=n_in)
        gpu_ctx.synchronize()
        w, g = want.cpu().numpy(), got.cpu().numpy()
    for b in range(blocks):
        assert util.count_bit_mismatch(g[b], w[b]) == 0, f"call {b}: max abs {np.abs(g[b] - w[b]).max()}"
    assert np.abs(w[1]).max() > 0.0 or np.abs(w[-1]).max() > 1e-6      # the FDN has produced output by the end
    assert np.abs(w[-1]).max() > 1e-6


def test_no_alloc_mono_input_and_accumulation(gpu_ctx):
    """mono input bypasses the scratch; the output is accumulated into like the reference's span"""
    import torch
    n, L, frames, blocks = 16, 12, 512, 10
    dev = torch.device("cuda", 0)
    stream = gpu_ctx.torch_stream()
    x = util.white_noise((blocks, n, frames), 5151).astype(np.float32)
    old, new = _banks(gpu_ctx, n, L)
    with torch.cuda.stream(stream):
        dx = torch.from_numpy(x).to(dev)
        want = torch.full((n, frames, L), 0.25, dtype=torch.float32, device=dev)
        got = torch.full((n, frames, L), 0.25, dtype=torch.float32, device=dev)
        gpu_ctx.synchronize()
        for b in range(blocks):
            old.ProcessInterleaved(dx[b], want, frames)
            new.ProcessInterleavedNoAlloc(dx[b], got, frames)
        gpu_ctx.synchronize()
        assert util.count_bit_mismatch(got.cpu().numpy(), want.cpu().numpy()) == 0


def test_no_alloc_rejects_sizes_beyond_the_static_scratch(gpu_ctx):
    import torch
    dev = torch.device("cuda", 0)
    frames, L = 480, 4
    n = J.REVERB_NO_ALLOC_MAX_FLOATS // frames + 1          # 513 instances x 480 frames: one row too many
    bank = J.ReverbBusBatch(gpu_ctx, n, FS, L, max_block_frames=frames)
    with torch.cuda.stream(gpu_ctx.torch_stream()):
        x = torch.zeros((n, frames, 2), dtype=torch.float32, device=dev)
        out = torch.zeros((n, frames, L), dtype=torch.float32, device=dev)
        gpu_ctx.synchronize()
        with pytest.raises(J.JplgError) as ei:
            bank.ProcessInterleavedNoAlloc(x, out, frames, in_channels=2)
        assert "exceeds the static scratch" in str(ei.value)
        bank.ProcessInterleavedNoAlloc(x[:, :, 0].contiguous(), out, frames)      # mono input of the same size: no scratch, accepted
        bank.ProcessInterleaved(x, out, frames, in_channels=2)                  # the allocating entry point takes the size
        gpu_ctx.synchronize()
