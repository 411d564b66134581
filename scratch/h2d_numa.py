"""H2D bandwidth of a 33.5 MB pinned block as a function of the NUMA node the pinned pages live on (first touch under a CPU affinity mask)."""
import glob, os, subprocess, torch
dev = torch.device("cuda", 0)
torch.cuda.init()
bdf = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", "0"], capture_output=True, text=True).stdout.strip().lower()
bdf = bdf[4:] if bdf.startswith("0000") and len(bdf) > 12 else bdf
try:
    print("gpu", bdf, "numa_node", open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
except Exception as e:
    print("numa_node unreadable", e)
subprocess.run("nvidia-smi topo -m | head -12", shell=True)
all_cpus = os.sched_getaffinity(0)
print("affinity", len(all_cpus), "cpus")


def parse(cl):
    out = set()
    for part in cl.strip().split(","):
        if "-" in part:
            a, b = part.split("-"); out |= set(range(int(a), int(b) + 1))
        elif part:
            out.add(int(part))
    return out


n = int(33.5e6) // 4
d = torch.empty(n, dtype=torch.float32, device=dev)
for node in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
    cpus = parse(open(node + "/cpulist").read()) & all_cpus
    if not cpus:
        print(os.path.basename(node), "no cpus allowed"); continue
    os.sched_setaffinity(0, cpus)
    h = torch.empty(n, dtype=torch.float32)
    h.fill_(1.0)                      # first touch on this node
    h = h.pin_memory()
    h2 = torch.ones(n, dtype=torch.float32).pin_memory()
    s = torch.cuda.Stream()
    for _ in range(3): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        e0.record(s)
        for _ in range(20): d.copy_(h, non_blocking=True)
        e1.record(s)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(os.path.basename(node), f"cpus {min(cpus)}..{max(cpus)} ({len(cpus)})  H2D 33.5 MB: {ms:.3f} ms = {n*4/ms/1e6:.1f} GB/s")
    os.sched_setaffinity(0, all_cpus)
    del h, h2
