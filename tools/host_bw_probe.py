"""dev tool (multi-rank): what limits the host-buffer path at 8 GPUs?  Prints the topology NVML reports for this rank's GPU
and measures concurrent pinned H2D + D2H bandwidth of all ranks at once, without and with CPU / memory binding.

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/host_bw_probe.py
"""
import ctypes
import os
import sys
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def log(*a):
    print(f"[rank {rank}]", *a, flush=True)


info = {}
try:
    import pynvml

    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(local)
    info["bus"] = pynvml.nvmlDeviceGetPciInfo(h).busId
    ncpu = os.cpu_count()
    words = (ncpu + 63) // 64
    try:
        aff = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * i + b for i, w in enumerate(aff) for b in range(64) if (w >> b) & 1]
        info["nvml_cpu_affinity"] = f"{len(cpus)} cpus {cpus[:4]}..{cpus[-4:]}" if cpus else "none"
    except Exception as e:
        cpus = []
        info["nvml_cpu_affinity"] = f"error {e}"
    try:
        mem = pynvml.nvmlDeviceGetMemoryAffinity(h, 4, 0)
        info["nvml_mem_affinity"] = [hex(x) for x in mem]
    except Exception as e:
        info["nvml_mem_affinity"] = f"error {e}"
except Exception as e:
    cpus = []
    info["nvml"] = f"error {e}"
info["sched_affinity"] = len(os.sched_getaffinity(0))
try:
    bus = info.get("bus", "")
    bus = bus.decode() if isinstance(bus, bytes) else bus
    dom, rest = bus.lower().split(":", 1)
    info["sysfs_numa_node"] = open(f"/sys/bus/pci/devices/{dom[-4:]}:{rest}/numa_node").read().strip()
except Exception as e:
    info["sysfs_numa_node"] = f"error {e}"
log(info)
if rank == 0:
    os.system("nvidia-smi topo -m 2>&1 | head -30; lscpu | grep -i -E 'numa|socket|model name|^CPU\\(s\\)'; ls /sys/devices/system/node/ 2>&1 | head; cat /proc/meminfo | head -3")


def measure(tag):
    n = 512 << 20
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_in.fill_(1)
    h_out.fill_(2)  # first touch under the current binding
    d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for mode in ("h2d", "d2h", "both"):
        for rep in range(2):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    d_a.copy_(h_in, non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    h_out.copy_(d_b, non_blocking=True)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        t = torch.tensor([dt], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        nbytes = n * (2 if mode == "both" else 1)
        res[mode] = round(world * nbytes / t.item() / 1e9, 1)
    if rank == 0:
        print(f"== {tag}: aggregate GB/s over {world} ranks: {res}", flush=True)


measure("no binding")
if cpus:
    os.sched_setaffinity(0, set(cpus) & os.sched_getaffinity(0) or os.sched_getaffinity(0))
    measure("bound to NVML cpu affinity (first touch on that node)")
if world > 1:
    dist.destroy_process_group()
