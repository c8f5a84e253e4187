"""Concurrent device-to-host bandwidth of all ranks (torchrun), with the page-locked destination allocated (a) wherever the
process happens to run and (b) after pinning the process to the CPUs NVML names as local to its GPU.  Diagnostic for the
sub-linear scaling of the host-buffer face (bench.py e2e at N = 4 / 8)."""
import ctypes as C
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from gym_usv_b200 import _native as nat  # noqa: E402

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
L = nat.lib()
N = 65536 * 143


def gpu_cpus():
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(local)
    bus = pynvml.nvmlDeviceGetPciInfo(h).busId
    bus = bus.decode() if isinstance(bus, bytes) else bus
    node = "?"
    try:
        node = open(f"/sys/bus/pci/devices/{bus[-12:].lower()}/numa_node").read().strip()
    except OSError:
        pass
    words = (os.cpu_count() + 63) // 64
    mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
    cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
    return bus, node, cpus


def pinned():
    p = C.c_void_p()
    nat.check(L.usvgpu_host_alloc(C.byref(p), N * 4))
    t = torch.frombuffer((C.c_float * N).from_address(p.value), dtype=torch.float32)
    t.zero_()
    return p, t


def bw(dst, src, reps=40):
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    g = torch.tensor([N * 4 * reps / dt / 1e9], device=dev, dtype=torch.float64)
    allg = [torch.zeros_like(g) for _ in range(world)]
    if world > 1:
        dist.all_gather(allg, g)
    else:
        allg = [g]
    return [float(x) for x in allg]


if rank == 0:
    for cmd in ("nvidia-smi topo -m", "lscpu | grep -i -E 'numa|socket|model name|^CPU\\(s\\)'", "numactl -H"):
        print("$", cmd, "\n" + subprocess.run(cmd, shell=True, capture_output=True, text=True).stdout, flush=True)
bus, node, cpus = gpu_cpus()
print(f"rank {rank}: GPU {local} bus {bus} numa_node {node} nvml cpu affinity {len(cpus)} cpus [{cpus[0] if cpus else '-'}..{cpus[-1] if cpus else '-'}]"
      f" process affinity {len(os.sched_getaffinity(0))} cpus", flush=True)
src = torch.zeros(N, device=dev)
p1, d1 = pinned()
assert d1.is_pinned(), "torch does not see the cudaHostAlloc'ed buffer as page-locked"
a = bw(d1, src)
if rank == 0:
    print("default placement: per-rank GB/s", ["%.1f" % x for x in a], "sum %.1f" % sum(a), flush=True)
if cpus:
    os.sched_setaffinity(0, cpus)
p2, d2 = pinned()
b = bw(d2, src)
if rank == 0:
    print("GPU-local CPUs   : per-rank GB/s", ["%.1f" % x for x in b], "sum %.1f" % sum(b), flush=True)
c = bw(d1, src)
if rank == 0:
    print("old buffers again: per-rank GB/s", ["%.1f" % x for x in c], "sum %.1f" % sum(c), flush=True)
if world > 1:
    dist.destroy_process_group()
