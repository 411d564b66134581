#!/usr/bin/env python3
"""bench.py — headline benchmark of the B200-native many-source spatial renderer.

Metric (BASELINE.json): real-time sources @ 48 kHz / 512-frame blocks -> 7.1.4.
Workload at every N: BASELINE config 3 per GPU — 16,384 mono sources, full block render
(MDAP gains -> Doppler fractional delay -> LR4 4-band filter -> gain-ramped mix to 7.1.4), weak scaling
(each rank owns 16,384 sources; the only collective is the reduce of the [512 x 12] speaker bus to rank 0).

One "step" = one 512-frame audio block for all sources of the job:
    (1) jplg_panner_mdap        new pan parameters for every source   -> mix maps
    (2) jplg_renderer_update    band gains, delays, mix maps
    (3) jplg_renderer_process_block   push + fused render + bus reduce
    (4) N > 1: torch.distributed.reduce(bus) over NCCL
value = sources * (block duration / step time) = how many sources this job renders in real time.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode parity|throughput] [--sources S]
  python bench.py --impl reference ...     times the reference's own CPU code (oracle/_ref) on the host cores
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FS = 48000.0
FRAMES = 512
L_OUT = 12
BLOCK_S = FRAMES / FS
SOURCES_PER_GPU = 16384
METRIC = "real-time sources @48kHz/512-frame->7.1.4 (full block render)"
UNIT = "sources"
# algorithmic bytes of the fused render kernel per mono source per 512-frame block, L = 12:
# ring read 512+2 samples, params (delay 4 + band gains 16 + mix map 48), state r/w 2*(32*4) + current mix r/w 2*48
RENDER_BYTES_PER_SOURCE = (FRAMES + 2) * 4 + 68 + 2 * 128 + 2 * 48
# whole step (all kernels): + input read 2048 + ring write 2176 (544 pushes, reference chunk quirk) + MDAP in/out 68
STEP_BYTES_PER_SOURCE = RENDER_BYTES_PER_SOURCE + FRAMES * 4 + 544 * 4 + 68


def log(*a):
    print(*a, file=sys.stderr, flush=True)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.thread = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.gpu), "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def count(self):
        return len(self.lines)

    def stop(self):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
            if self.thread:
                self.thread.join(timeout=2)
        sm, mx, reasons = [], 0.0, set()
        for line in self.lines:
            s = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(s[1])); mx = max(mx, float(s[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm),
                "window": "warm-up + timed region + e2e region (100 ms period)"}


def make_inputs(n_sources, pool, seed):
    """Synthetic scene: white noise per source, linear trajectories (1-300 m, +-30 m/s), random band gains,
    random pan parameters per block. `pool` distinct blocks that the steps cycle through (pool * 33.5 MB > L2)."""
    from tests import util
    audio = util.white_noise((pool, n_sources, FRAMES, 1), seed + 1)
    params = np.stack([util.pan_params(n_sources, seed + 10 + k, edge=False) for k in range(pool)])
    band = (0.1 + 0.9 * util.uniform01(pool * n_sources * 4, seed + 2)).reshape(pool, n_sources, 4).astype(np.float32)
    dist0 = 1.0 + 299.0 * util.uniform01(n_sources, seed + 3)
    speed = (util.uniform01(n_sources, seed + 4) * 2.0 - 1.0) * 30.0
    t = (np.arange(pool, dtype=np.float32) * np.float32(BLOCK_S))[:, None]
    delay = (np.maximum(dist0[None] + speed[None] * t, 0.5) / np.float32(343.2)).astype(np.float32)
    return audio, params, band, delay


# ------------------------------------------------------------------------------------------------
def run_reference_arm(args, rank, world):
    """Times the reference's own CPU implementation of the path (oracle/_ref: the unmodified reference compiled from
    /root/reference) on the host cores, all hardware threads, on a bounded sample of the same workload."""
    if rank != 0:
        return
    from oracle import refpy
    kind = "reference"
    if not refpy.available("sse2"):
        _emit(({"impl": "reference", "unavailable": "oracle/_ref/libjplref_sse2.so not present (reference tree absent at build time)"}))
        return
    cores = os.cpu_count() or 1
    n = args.ref_sources
    ref = refpy.Ref("sse2")
    M = __import__("jplspatial_b200").ChannelMask
    audio, params, band, delay = make_inputs(n, 4, 0x5EED)
    rp = ref.panner(M.Surround_7_1_4)
    rl = rp.layout(M.Mono)
    mix0 = rp.mdap(rl, params[0]).reshape(n, L_OUT)
    fx = [ref.dse(FS, 1, L_OUT, band[0, i], float(delay[0, i]), mix0[i]) for i in range(n)]
    bus = np.zeros((FRAMES, L_OUT), np.float32)

    def step(k):
        j = k % audio.shape[0]
        mix = rp.mdap(rl, params[j], threads=cores).reshape(n, L_OUT)
        ref.dse_batch_update(fx, band[j], delay[j], mix)
        bus[:] = 0
        ref.dse_batch_process(fx, audio[j].reshape(n, -1), FRAMES, bus.reshape(-1), threads=cores)

    for k in range(args.warmup):
        step(k)
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(args.warmup + k)
    dt = (time.perf_counter() - t0) / args.steps
    value = n * BLOCK_S / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "BASELINE config 3: full block render, mono sources -> 7.1.4, 48 kHz / 512 frames (MDAP gains + update + render per block)",
                   "sources_per_step": n, "note": "bounded sample of the 16,384-source workload; value = sample sources x block duration / step time"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{n} mono sources x {args.steps} blocks of 512 frames, {cores} std::threads, reference SSE2 build (-O2)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    _emit(line)


def cpu_baseline_sample(log_fn):
    """Bounded CPU sample for the `cpu_baseline` object of the GPU arm (rank 0, N = 1)."""
    from oracle import refpy
    cores = os.cpu_count() or 1
    if refpy.available("sse2"):
        ref = refpy.Ref("sse2")
        kind = "reference"
    else:
        return None
    import jplspatial_b200 as J
    n, blocks = 1024, 24
    audio, params, band, delay = make_inputs(n, 4, 0xC0FFEE)
    rp = ref.panner(J.ChannelMask.Surround_7_1_4)
    rl = rp.layout(J.ChannelMask.Mono)
    mix0 = rp.mdap(rl, params[0]).reshape(n, L_OUT)
    fx = [ref.dse(FS, 1, L_OUT, band[0, i], float(delay[0, i]), mix0[i]) for i in range(n)]
    bus = np.zeros((FRAMES, L_OUT), np.float32)
    t0 = time.perf_counter()
    for k in range(blocks):
        j = k % 4
        mix = rp.mdap(rl, params[j], threads=cores).reshape(n, L_OUT)
        ref.dse_batch_update(fx, band[j], delay[j], mix)
        ref.dse_batch_process(fx, audio[j].reshape(n, -1), FRAMES, bus.reshape(-1), threads=cores)
    dt = (time.perf_counter() - t0) / blocks
    # single-thread figure for context
    t1 = time.perf_counter()
    for k in range(4):
        ref.dse_batch_process(fx[:256], audio[k % 4, :256].reshape(256, -1), FRAMES, bus.reshape(-1), threads=1)
    us1 = (time.perf_counter() - t1) / (4 * 256) * 1e6
    log_fn(f"cpu baseline: {dt * 1e3:.2f} ms per {n}-source block on {cores} threads; single-thread DirectSoundEffect {us1:.1f} us/source-block")
    return {"value": n * BLOCK_S / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{n} mono sources x {blocks} blocks of 512 frames (MDAP + update + render), {cores} std::threads, reference SSE2 build; "
                      f"1-thread DirectSoundEffect = {us1:.1f} us per source-block"}


# ------------------------------------------------------------------------------------------------
_JSON_OUT = None


def bind_to_gpu_numa_node(torch, local_rank):
    """Run this rank (and first-touch its pinned buffers) on the NUMA node its GPU hangs off: with one rank per GPU on a two-socket box the host side of
    eight concurrent 33.5 MB-per-block input streams otherwise crosses the socket interconnect. Returns the node, or None when the box does not expose one."""
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus |= set(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception as e:          # purely an optimisation
        log(f"[bench] NUMA binding skipped: {e}")
        return None


def _claim_stdout():
    """The driver reads ONE JSON line from stdout: keep a private handle to the real stdout for it and point fd 1 at stderr for everything else
    (NCCL prints its version banner to stdout on the rank that initialises first)."""
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def _emit(line):
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="parity", choices=["parity", "throughput"])
    ap.add_argument("--sources", type=int, default=SOURCES_PER_GPU, help="sources per GPU")
    ap.add_argument("--ring", type=int, default=0, help="ring size per source (0 = reference value bit_ceil(2*fs) = 131072)")
    ap.add_argument("--pool", type=int, default=8, help="distinct input blocks cycled through (pool x 33.5 MB > L2)")
    ap.add_argument("--ref-sources", type=int, default=2048)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-push-ahead", action="store_true",
                    help="issue every block's ring pushes inside process_block instead of one block ahead on a side stream (jplg_renderer_push_ahead)")
    ap.add_argument("--no-pan-overlap", action="store_true",
                    help="compute the panning data on the rendering context's stream instead of a background context (jplg_context_create_ex) "
                         "that works one block ahead on the SMs the render leaves free")
    ap.add_argument("--no-fused-update", action="store_true",
                    help="jplg_panner_mdap + jplg_renderer_update as two launches with the mix maps in between, instead of jplg_renderer_update_panned")
    ap.add_argument("--bus", choices=["peer", "nccl", "none"], default="peer",
                    help="multi-GPU combine of the partial buses: peer = jplg_bus_exchange (stores into the root GPU's memory over NVLink, rank-ordered sum), "
                         "nccl = torch.distributed reduce, double-buffered")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import jplspatial_b200 as J

    if not torch.cuda.is_available() or J.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device — this framework has no CPU fallback (use --impl reference for the CPU reference arm)")
    torch.cuda.set_device(local_rank)
    numa_node = bind_to_gpu_numa_node(torch, local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    N = args.sources
    mode = J.MODE_PARITY if args.mode == "parity" else J.MODE_THROUGHPUT
    ctx = J.Context(local_rank)
    stream = ctx.torch_stream()
    audio, params, band, delay = make_inputs(N, args.pool, 0xB200 + rank)
    pan = J.VBAPanner(ctx)
    assert pan.Initialize(J.ChannelMask.Surround_7_1_4)
    lay = pan.InitializeSourceLayout(J.ChannelMask.Mono)
    mix0 = pan.ProcessVBAPData(lay, params[0]).reshape(N, L_OUT)
    rend = J.DirectSoundBatch(ctx, FS, N, 1, L_OUT, band[0], delay[0], mix0, ring_size=args.ring, max_block_frames=FRAMES, mode=mode)
    dev = torch.device("cuda", local_rank)

    with torch.cuda.stream(stream):
        d_audio = torch.from_numpy(audio).to(dev)
        d_params = torch.from_numpy(params).to(dev)
        d_band = torch.from_numpy(band).to(dev)
        d_delay = torch.from_numpy(delay).to(dev)
        d_mix = torch.empty((N, 1, L_OUT), dtype=torch.float32, device=dev)
    ctx.synchronize()

    from jplspatial_b200.sharding import PipelinedBusReduce, make_peer_bus_exchange
    # the only exchange of the path: the sum of the per-GPU partial buses ([512][12] floats per rank and block)
    use_peer = dist is not None and args.bus == "peer"
    # nccl mode: two bus buffers, the NCCL reduce of block k overlaps the render of block k+1
    reducer = PipelinedBusReduce(FRAMES, L_OUT, dev, depth=2, dst=0)
    if args.bus == "none":
        reducer.submit = lambda k: None     # diagnostic only: N independent replicas, no combine (the line's `config.parallelism` says so)
    exchange = make_peer_bus_exchange(ctx, FRAMES * L_OUT) if use_peer else None
    with torch.cuda.stream(stream):
        local_bus = torch.zeros((FRAMES, L_OUT), dtype=torch.float32, device=dev)
        final_bus = torch.zeros((FRAMES, L_OUT), dtype=torch.float32, device=dev)
    seq = [0]

    push_ahead = not args.no_push_ahead
    gk = [0]
    # game-thread / audio-thread split of the reference (SpatialManager::AdvanceSimulation beside the audio callback): block k+1's panning data is
    # computed by a background context while block k renders; its stream has the lower priority, so k_render's 128 CTAs are placed first and
    # k_mdap_3d fills the remaining SMs. Two mix-map buffers; events order the two streams.
    pan_overlap = not args.no_pan_overlap
    fused_update = not args.no_fused_update       # panning kernel writes the renderer's next target set itself (jplg_renderer_update_panned)
    if pan_overlap:
        ctx_pan = J.Context(local_rank, background=True)
        stream_pan = ctx_pan.torch_stream()
        pan_bg = J.VBAPanner(ctx_pan)
        assert pan_bg.Initialize(J.ChannelMask.Surround_7_1_4)
        lay_bg = pan_bg.InitializeSourceLayout(J.ChannelMask.Mono)
        with torch.cuda.stream(stream):
            d_mix_bg = [torch.empty((N, 1, L_OUT), dtype=torch.float32, device=dev) for _ in range(2)]
        ctx.synchronize()
        ev_mdap = [torch.cuda.Event() for _ in range(2)]
        ev_upd = torch.cuda.Event()
        mdap_issued = [-1]

        def issue_mdap(k, after_update):
            # (the library enqueues on its contexts' own streams whatever torch's current stream is: no stream context managers in the step)
            j = k % args.pool
            if after_update:
                if not fused_update:
                    stream_pan.wait_event(ev_upd)  # the buffer's previous reader (update k-1) is older than this event
                rend.WaitRenderStarted(ctx_pan)    # only once block k-1's render is resident: k_mdap_3d takes the SMs it leaves free
            if fused_update:
                rend.UpdatePanned(pan_bg, lay_bg, v_params[j], v_band[j], v_delay[j])      # the library orders it against the renders itself
            else:
                pan_bg.ProcessVBAPData(lay_bg, v_params[j], out=d_mix_bg[k & 1])
            ev_mdap[k & 1].record(stream_pan)
            mdap_issued[0] = k

        def join_pan():
            if mdap_issued[0] >= 0:
                stream.wait_event(ev_mdap[mdap_issued[0] & 1])

    # per-block views made once: the step itself is host-side plumbing and has to stay well under the 0.14 ms the device needs per block
    v_params = [d_params[j] for j in range(args.pool)]
    v_band = [d_band[j] for j in range(args.pool)]
    v_delay = [d_delay[j] for j in range(args.pool)]
    v_audio = [d_audio[j] for j in range(args.pool)]
    bus_ptr, bus_floats = rend.device_bus_ptr(), FRAMES * L_OUT
    peer_out = final_bus if rank == 0 else None

    def step_device(_k):
        k = gk[0]          # consecutive block numbers whatever the caller's phase, so the block pushed ahead is always the next one rendered
        gk[0] += 1
        j = k % args.pool
        if pan_overlap:
            if mdap_issued[0] != k:
                issue_mdap(k, False)
            if not fused_update:
                stream.wait_event(ev_mdap[k & 1])
                rend.UpdateParameters(v_band[j], v_delay[j], d_mix_bg[k & 1])
                ev_upd.record(stream)
        elif fused_update:
            rend.UpdatePanned(pan, lay, v_params[j], v_band[j], v_delay[j])
        else:
            pan.ProcessVBAPData(lay, v_params[j], out=d_mix)
            rend.UpdateParameters(v_band[j], v_delay[j], d_mix)
        if use_peer:
            # every rank's k_bus_reduce result goes straight into the root GPU's gather buffer (peer stores over NVLink); the root adds the
            # slots in rank order. No collective library on the data path.
            rend.ProcessInterleaved(v_audio[j], local_bus, FRAMES)
            exchange.submit(bus_ptr, bus_floats, seq[0], out=peer_out)
            seq[0] += 1
        else:
            with torch.cuda.stream(stream):       # torch ops (the bus zero-fill, NCCL's stream ordering) follow torch's current stream
                bus = reducer.begin(k)
                rend.ProcessInterleaved(v_audio[j], bus, FRAMES)
                reducer.submit(k)
        if push_ahead:
            # throughput rendering: block k+1's input is resident, so its ring pushes (an HBM-bound copy) are issued right after block k's render
            # has been enqueued and run beside it on the SMs the render leaves free; block k only waits for the pushes issued one step earlier
            rend.PushAhead(v_audio[(k + 1) % args.pool], FRAMES)
        if pan_overlap:
            issue_mdap(k + 1, True)

    def barrier():
        with torch.cuda.stream(stream):
            if pan_overlap:
                join_pan()
            reducer.drain()
            if exchange is not None:
                exchange.wait()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (`value`)
    sampler = ClockSampler(local_rank)
    sampler.start()
    for k in range(args.warmup):
        step_device(k)
    barrier()
    # only the dominant kernel is bracketed by events inside the timed region (two event records per step); the per-kernel
    # breakdown of the other kernels is taken in a separate short pass afterwards so its events do not stretch the timed steps
    ctx.set_kernel_timing(True, only=("render",))
    ctx.reset_kernel_timing()
    launches0 = ctx.launch_count + (ctx_pan.launch_count if pan_overlap else 0)
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    barrier()
    with torch.cuda.stream(stream):
        ev0.record(stream)
    t_host0 = time.perf_counter()
    for k in range(args.steps):
        step_device(args.warmup + k)
    host_ms_step = (time.perf_counter() - t_host0) * 1e3 / args.steps     # host time to enqueue a step: must stay below the device's step time
    with torch.cuda.stream(stream):
        if pan_overlap:
            join_pan()               # the panning data computed one block ahead by the last step is inside the timed region too
        reducer.drain()              # the last blocks' reduces are inside the timed region
        if exchange is not None:
            exchange.wait()          # ... and so is the peer-memory combine of the last blocks
        ev1.record(stream)
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    # fingerprint of the last timed block's (combined) bus: the same for every pipelining option at equal --warmup / --steps / --pool / --gpus
    import zlib
    last_bus = final_bus if use_peer else reducer.result(gk[0] - 1)
    bus_crc = zlib.crc32(last_bus.cpu().numpy().tobytes()) if rank == 0 else 0
    launches = ctx.launch_count + (ctx_pan.launch_count if pan_overlap else 0) - launches0  # our kernels only (torch's bus zero-fill and NCCL are not counted)
    ktimes = {"render": ctx.kernel_time("render")}
    ctx.set_kernel_timing(False)
    bsteps = max(10, min(args.steps, 40))
    ctx.set_kernel_timing(True)
    ctx.reset_kernel_timing()
    if pan_overlap:
        ctx_pan.set_kernel_timing(True)
        ctx_pan.reset_kernel_timing()
    for k in range(bsteps):
        step_device(args.warmup + args.steps + k)
    barrier()
    breakdown = {name: ctx.kernel_time(name) for name in ("mdap", "update", "push", "render", "bus_reduce")}
    ctx.set_kernel_timing(False)
    if pan_overlap:
        # measured while sharing the device with k_render (and the pushes): a duration under overlap, not the kernel's own cost
        breakdown["mdap"] = ctx_pan.kernel_time("mdap")
        ctx_pan.set_kernel_timing(False)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    per_rank_ms = [ms_total / args.steps]
    if dist is not None:
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        per_rank_ms = [float(x.item()) / args.steps for x in allt]
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = N * world * BLOCK_S / (ms_step * 1e-3)

    # ---- end to end through the C-ABI with HOST (pinned) buffers: H2D of the step's inputs and D2H of the bus inside the timed region
    e2e = None
    if not args.no_e2e:
        def pinned(a):
            t_ = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
            return t_, t_.numpy()
        keep = []
        h_audio, h_params, h_band, h_delay = [], [], [], []
        for j in range(args.pool):
            for src, dst in ((audio[j], h_audio), (params[j], h_params), (band[j], h_band), (delay[j], h_delay)):
                tt, nn = pinned(src); keep.append(tt); dst.append(nn)
        tt, h_mix = pinned(np.zeros((N, 1, L_OUT), np.float32)); keep.append(tt)
        tt, h_bus = pinned(np.zeros((FRAMES, L_OUT), np.float32)); keep.append(tt)
        d_partial = torch.zeros((FRAMES, L_OUT), dtype=torch.float32, device=dev)

        # two bus slots: block k is submitted (JPLG_MEM_HOST_ASYNC: nothing waits on the host, the 33.5 MB input transfer runs on the library's copy
        # stream while block k-1 renders) and then block k-1 is collected; every step still moves its own inputs H2D and its own bus D2H
        tb, h_bus2 = pinned(np.zeros((FRAMES, L_OUT), np.float32)); keep.append(tb)
        h_buses = (h_bus, h_bus2)
        if use_peer:
            with torch.cuda.stream(stream):
                d_final2 = [torch.zeros((FRAMES, L_OUT), dtype=torch.float32, device=dev) for _ in range(2)]
            t_final, h_final = [], []
            for _ in range(2):
                tt, nn = pinned(np.zeros((FRAMES, L_OUT), np.float32)); t_final.append(tt); h_final.append(nn)
            ev_final = [torch.cuda.Event() for _ in range(2)]
            ev_read = [torch.cuda.Event() for _ in range(2)]
            rd_stream = torch.cuda.Stream(device=dev)
            ctx.synchronize()

        def submit_host(k):
            j = k % args.pool
            hb = h_buses[k & 1]
            pan.ProcessVBAPData(lay, h_params[j], out=h_mix, async_host=True)          # H2D params, D2H mix maps (enqueued)
            rend.UpdateParameters(h_band[j], h_delay[j], h_mix, async_host=True)        # H2D
            hb[:] = 0.0
            rend.ProcessInterleaved(h_audio[j], hb, FRAMES, async_host=True)            # H2D audio (copy stream) + bus, D2H bus
            if use_peer:
                # N > 1: the partial buses are combined on the device (peer-memory exchange, as in the device-resident pass) and the root reads the
                # combined bus back; the other ranks' own partial buses still travel D2H with their blocks
                exchange.submit(bus_ptr, bus_floats, seq[0], out=d_final2[k & 1] if rank == 0 else None)
                seq[0] += 1
                if rank == 0:
                    exchange.wait()
                    ev_final[k & 1].record(stream)
                    # read-back on a torch-owned stream (torch's pinned-memory allocator remembers the streams a pinned block was used on, and the
                    # library's stream dies with its context)
                    rd_stream.wait_event(ev_final[k & 1])
                    with torch.cuda.stream(rd_stream):
                        t_final[k & 1].copy_(d_final2[k & 1], non_blocking=True)
                        ev_read[k & 1].record(rd_stream)

        def collect_host(k):
            rend.WaitBlock()
            hb = h_buses[k & 1]
            if use_peer:
                if rank == 0:
                    ev_read[k & 1].synchronize()
                    return float(h_final[k & 1][0, 0])
            elif dist is not None:
                d_partial.copy_(torch.from_numpy(hb))
                dist.reduce(d_partial, dst=0)
                if rank == 0:
                    hb[:] = d_partial.cpu().numpy()
            return float(hb[0, 0])

        e2e_steps = max(10, min(args.steps, 100))
        # warm-up: the device-resident pass above left the PCIe link idle for ~a second and the first tens of transfers after that run at a fraction
        # of the link's rate (measured on these boxes: 16-40 GB/s for the first copies, 50-55 GB/s once it has ramped up)
        e2e_warm = max(args.warmup, 30)
        for k in range(e2e_warm):
            submit_host(k)
            if k:
                collect_host(k - 1)
        collect_host(e2e_warm - 1)
        barrier()
        t0 = time.perf_counter()
        for k in range(e2e_steps):
            submit_host(e2e_warm + k)
            if k:
                collect_host(e2e_warm + k - 1)
        collect_host(e2e_warm + e2e_steps - 1)
        barrier()
        dt = (time.perf_counter() - t0) / e2e_steps
        tt2 = torch.tensor([dt], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tt2, op=dist.ReduceOp.MAX)
        dt = float(tt2.item())
        h2d = N * FRAMES * 4 + N * 20 + N * 16 + N * 4 + N * L_OUT * 4 + FRAMES * L_OUT * 4
        d2h = N * L_OUT * 4 + FRAMES * L_OUT * 4 + (FRAMES * L_OUT * 4 if use_peer else 0)     # per rank; the last term is the root's combined bus
        e2e = {"value": N * world * BLOCK_S / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "path": "C ABI with pinned HOST buffers, JPLG_MEM_HOST_ASYNC: block k+1's input transfer overlaps block k's render (2 blocks in flight); the audio goes through the copy engine, the small arrays and the bus are read / written by the kernels through the buffers' device alias (same bytes over PCIe)"
                       + ("; N > 1: partial buses combined on the device by the peer-memory exchange, the root reads the combined bus back" if use_peer else "")}

    # a short run (--no-e2e, few steps) can end before nvidia-smi has delivered its first line: keep the GPU under the same load until it has
    # (every rank takes the same decision: the steps contain the bus exchange)
    extra, t_end = 0, time.time() + 3.0
    while True:
        need = 1 if (sampler.proc is not None and sampler.count() == 0 and time.time() < t_end) else 0
        if dist is not None:
            tn = torch.tensor([need], dtype=torch.int32, device=dev)
            dist.all_reduce(tn, op=dist.ReduceOp.MAX)
            need = int(tn.item())
        if not need:
            break
        for k in range(64):
            step_device(extra + k)
        extra += 64
        barrier()
    clocks = sampler.stop()
    if exchange is not None and exchange.error() != 0:
        raise SystemExit(f"bench.py: bus exchange reported a device-side wait timeout (error {exchange.error()}) on rank {rank}")
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (k_render), measured live with CUDA events on the launching stream
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak = float(json.load(open(peaks_path))["hbm_gbs"]); peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak = 6650.0; peak_src = "fallback (B200_PROFILING.md)"
    r_ms, r_n = ktimes["render"]
    r_avg_ms = r_ms / max(r_n, 1)
    achieved = RENDER_BYTES_PER_SOURCE * N / (r_avg_ms * 1e-3) / 1e9
    ncu_traffic = None
    tpath = os.path.join(ROOT, "profiles", "render_traffic.json")
    if os.path.exists(tpath):
        try:
            ncu_traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            pass
    roofline = {"bound": "hbm", "kernel": "k_render", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic, "peak_source": peak_src, "avg_launch_ms": r_avg_ms,
                "algorithmic_bytes_per_launch": RENDER_BYTES_PER_SOURCE * N,
                "whole_step_gbs": STEP_BYTES_PER_SOURCE * N / (ms_step * 1e-3) / 1e9,
                "note": "fused render is FP32-issue bound (~200 FP32 ops/sample in parity mode), not HBM bound: see DESIGN.md §5",
                "kernel_ms_per_step": {k: v[0] / max(bsteps, 1) for k, v in breakdown.items()},
                "kernel_ms_per_step_note": f"separate pass of {bsteps} steps with every kernel bracketed by events (not part of the timed region)"}

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        try:
            cpu = cpu_baseline_sample(log)
        except Exception as e:  # the checker is optional at bench time
            log("cpu baseline failed:", e)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "BASELINE config 3: full block render, 16,384 mono sources per GPU -> 7.1.4, 48 kHz / 512 frames "
                               "(MDAP gains + update + push + fused delay/LR4/mix + bus reduce per block)",
                   "sources_per_gpu": N, "total_sources": N * world, "mode": args.mode, "ring_size": rend.ring_size,
                   "host_numa_node": numa_node,
                   "pipelining": {"push_ahead": push_ahead, "pan_overlap": pan_overlap, "fused_update": fused_update,
                                  "note": "every step runs all five kernels once; with push_ahead / pan_overlap block k+1's ring pushes / panning data "
                                          "are produced on lower-priority streams while block k renders"},
                   "l2_policy": f"inputs larger than L2: {args.pool} distinct 33.5 MB input blocks cycled, {N * rend.ring_size * 4 / 2**30:.1f} GiB of rings",
                   "parallelism": (f"sources sharded over {world} GPU(s); partial [512x12] buses stored into the root GPU's gather buffer over NVLink peer memory and summed in rank order (jplg_bus_exchange, no collective library on the data path)"
                                   if use_peer else "DIAGNOSTIC: independent replicas, buses NOT combined" if args.bus == "none" else f"sources sharded over {world} GPU(s); NCCL reduce of the [512x12] bus, double-buffered (block k reduce overlaps block k+1 render)") if world > 1 else "single GPU"},
        "per_rank_ms_per_step": [round(x, 5) for x in per_rank_ms], "host_enqueue_ms_per_step": round(host_ms_step, 5), "bus_crc32": f"{bus_crc:08x}", "clocks": clocks, "gpu_launches": launches, "e2e": e2e, "roofline": roofline, "cpu_baseline": cpu,
    }
    _emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
