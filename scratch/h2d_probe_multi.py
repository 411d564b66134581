"""Bare pinned host-to-device probe, one rank per GPU (torchrun): every rank copies 33.5 MB per iteration (config 3's audio per block) from
pinned host memory, no kernels. Variants: plain pinned, write-combined pinned, 4 chunks on 2 copy streams. Prints per-rank and aggregate GB/s:
the ceiling bench.py's e2e pass can reach on this box at N ranks."""
import ctypes, json, os, sys, time
import torch
import torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1: dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
nbytes = 16384 * 512 * 4
n = nbytes // 4
cudart = ctypes.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else None
def pinned(wc):
    if not wc or cudart is None: return torch.empty(n, dtype=torch.float32).pin_memory(), None
    p = ctypes.c_void_p()
    cudart.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
    assert cudart.cudaHostAlloc(ctypes.byref(p), nbytes, 4) == 0      # cudaHostAllocWriteCombined
    return None, p
res = {}
d = torch.empty(n, dtype=torch.float32, device=dev)
streams = [torch.cuda.Stream(), torch.cuda.Stream()]
for variant in ("pinned", "pinned_4chunks_2streams", "write_combined"):
    h, raw = pinned(variant == "write_combined")
    def step():
        if raw is not None:
            cudart.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
            cudart.cudaMemcpyAsync(ctypes.c_void_p(d.data_ptr()), raw, nbytes, 1, ctypes.c_void_p(streams[0].cuda_stream))
        elif variant == "pinned":
            with torch.cuda.stream(streams[0]): d.copy_(h, non_blocking=True)
        else:
            q = n // 4
            for c in range(4):
                with torch.cuda.stream(streams[c & 1]): d[c * q:(c + 1) * q].copy_(h[c * q:(c + 1) * q], non_blocking=True)
    for _ in range(40): step()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    t0 = time.perf_counter()
    iters = 100
    for _ in range(iters): step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / iters
    t = torch.tensor([dt], device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[variant] = {"ms_per_copy_max_over_ranks": round(float(t.item()) * 1e3, 4), "per_rank_gbs": round(nbytes / float(t.item()) / 1e9, 1),
                    "aggregate_gbs": round(world * nbytes / float(t.item()) / 1e9, 1)}
    if raw is not None: cudart.cudaFreeHost(raw)
if rank == 0: print(json.dumps({"ranks": world, "bytes_per_rank_per_copy": nbytes, "host_cpus": os.cpu_count(), **res}))
if world > 1: dist.destroy_process_group()
