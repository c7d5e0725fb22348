#!/usr/bin/env python
"""bench.py — the hot path of BASELINE.json on N B200s of one node.

Metric: Mrays/s nearest-hit at 1M AABBs (+ BVH refit+rebuild ms, + % of HBM roofline).
Workload (configs[2]): 1M uniform AABBs in [-1000,1000]^3, incoherent random rays; weak scaling —
every rank traces `--rays` rays (default 16M) against the replicated BVH.  A step = one pass of the
batched GetRaycast over the rank's ray batch.

  python bench.py --gpus 1 --steps 5 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
  python bench.py --impl reference ...      # the reference's own CPU GetRaycast on the host cores

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

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

METRIC = "Mrays/s nearest-hit at 1M AABBs"
UNIT = "Mrays/s"


def env_int(k, d):
    return int(os.environ.get(k, d))


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def make_scene(n_colliders: int, workload: str = "config3"):
    from engine_b200 import synth

    if workload == "config5":  # clustered / skewed: Gaussian clusters + 5 % uniform background
        return synth.clustered_boxes(n_colliders, seed=5)
    return synth.uniform_boxes(n_colliders, seed=3, extent=1000.0, hmin=0.5, hmax=2.0)


def make_rays(n: int, rank: int, workload: str = "config3"):
    from engine_b200 import synth

    return synth.random_rays(n, seed=4 + 1000 * rank, extent=2000.0 if workload == "config5" else 1000.0)


# -------------------------------------------------------------------------------------------------
# CPU arm: the reference's own GetRaycast (oracle/_ref) — or the oracle port when the reference
# library did not travel — on the host cores, on a bounded sample of the same workload.
# -------------------------------------------------------------------------------------------------
class CpuArm:
    def __init__(self, scene, want_threads: int | None = None):
        from oracle.oracle import Port, RefEngine

        self.port = Port()
        self.threads = want_threads or (os.cpu_count() or 1)
        t0 = time.time()
        if RefEngine.available():
            self.kind = "reference"
            self.ref = RefEngine()
            self.ref.create_entities(scene.parent[1:], scene.pos[1:], scene.rot[1:], scene.scl[1:])
            self.ref.add_colliders(scene.collider_entity, scene.collider_aabb)
            self.ref.transform_update()  # the reference's TransformSystem::onUpdate
            ent, box, _ = self.ref.refit(scene.n_colliders)  # the reference's transformBox loop
            # BuildNode is O(N^2) (est. 53 h at 1M): the tree comes from the oracle's binned-SAH builder,
            # which minimises the same cost, and is assigned to the reference's public bvh_ member
            self.tree = self.port.build_binned_sah(ent, box)
            self.ref.set_bvh(self.tree)
        else:
            self.kind = "port"
            self.ref = None
            world = self.port.transform_update(scene.parent, scene.pos, scene.rot, scene.scl)
            box = self.port.refit(scene.collider_entity, scene.collider_aabb, world)
            self.tree = self.port.build_binned_sah(scene.collider_entity, box)
        self.setup_s = time.time() - t0

    def trace(self, rays) -> float:
        """seconds for one pass over `rays` with all host threads"""
        if self.ref is not None:
            _, secs = self.ref.raycast(rays, nthreads=self.threads, want_secs=True)
            return secs
        t0 = time.time()
        self.port.raycast_tree(self.tree, rays, nthreads=self.threads)
        return time.time() - t0

    def calibrate(self, rays_pool, target_s: float) -> int:
        probe = min(len(rays_pool), 20000)
        s = max(self.trace(rays_pool[:probe]), 1e-4)
        n = int(probe / s * target_s)
        return int(min(max(n, 20000), len(rays_pool)))


def run_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    scene = make_scene(args.colliders, args.workload)
    arm = CpuArm(scene)
    pool = make_rays(min(args.rays, 4_000_000), 0, args.workload)
    n = arm.calibrate(pool, target_s=3.0)
    rays = pool[:n]
    for _ in range(args.warmup):
        arm.trace(rays)
    secs = [arm.trace(rays) for _ in range(args.steps)]
    total = float(sum(secs))
    v = n * args.steps / total / 1e6
    sample = f"{n} of the workload's rays per step ({arm.kind} GetRaycast, binned-SAH tree injected into bvh_; the reference's O(N^2) BuildNode cannot build 1M)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(args),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": arm.threads, "kind": arm.kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "setup_s": arm.setup_s,
    }))


def workload_config(args):
    if args.workload == "config5":
        return {
            "workload": "configs[4]: synthetic clustered AABBs (1000 Gaussian clusters, Zipf populations, 5 % uniform background, seed 5), incoherent random rays (seed 4), T=1000",
            "colliders": args.colliders, "rays_per_gpu": args.rays, "sharding": "rays sharded across ranks, BVH replicated (NCCL broadcast)",
            "l2": "ray + hit streams exceed the 126 MB L2 every step; at 10M colliders the 640 MB BVH does too",
        }
    return {
        "workload": "configs[2]: synthetic 1M AABBs (uniform in [-1000,1000]^3, half-extents U[0.5,2], seed 3), incoherent random rays (seed 4), T=1000",
        "colliders": args.colliders, "rays_per_gpu": args.rays, "sharding": "rays sharded across ranks, BVH replicated (NCCL broadcast)",
        "l2": "ray + hit streams (60 B/ray, 960 MB per step at 16M rays) exceed the 126 MB L2 every step; the 64 MB BVH is legitimately L2-resident",
    }


# -------------------------------------------------------------------------------------------------
class DevMem:
    """device memory owned by the library (the gather window) as a torch tensor: torch.as_tensor(DevMem(addr, n), device=...)"""

    def __init__(self, addr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (addr, False), "version": 2}


def source_hash() -> str:
    """identifies the build a committed ncu capture belongs to (roofline.traffic): sha256 over the sources k_trace_lbvh is compiled from"""
    import hashlib

    h = hashlib.sha256()
    d = os.path.join(ROOT, "engine_b200", "csrc")
    for f in ("b2col_trace.inl", "b2col_bvh.cuh", "b2col_math.cuh"):
        h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def measure_rebuild(cs, scene, args, torch):
    """refit + rebuild per frame at N colliders (BASELINE.json metric, second half), three ways:
      ms          every collider MOVES every frame: a device op perturbs all positions, the TRS arrays are handed over with
                  b2col_set_trs_device (44 B/entity device-to-device) and b2col_update rebuilds; ms = the update alone, frame_ms =
                  the whole frame incl. animation and hand-over — medians of 100
      static_ms   the same TRS replayed (nothing moves, warmest caches): the round-1 figure, kept for continuity
      refit_only_ms  b2col_refit on the moving scene (same tree shape, new boxes)"""
    N = scene.n_colliders
    peak, _ = measured_peak_gbs()
    algo = 280.0 * N  # refit 140 B + build 140 B per collider
    cs.set_scene(scene)
    for _ in range(max(args.warmup, 3) + 2):
        cs.on_update()  # synchronous: the key-width and Morton-plan policies read what the previous builds reported and settle within four frames
    K = max(args.steps, 10)
    l0 = cs.stats()["kernel_launches"]
    cs.event_record(0)
    for _ in range(K):
        cs.on_update(wait=False)
    cs.event_record(1)
    static_ms = cs.event_elapsed_ms(0, 1) / K
    launches_per_update = (cs.stats()["kernel_launches"] - l0) // K
    # moving scene: everything on the context's own stream (torch ops included), no host synchronisation inside the loop
    ext = torch.cuda.ExternalStream(cs.stream)
    d_pos, d_rot, d_scl = (torch.from_numpy(a).cuda() for a in (scene.pos, scene.rot, scene.scl))
    g = torch.Generator(device="cuda").manual_seed(7)
    frames = 100
    delta = 0.05 * torch.randn((8, scene.pos.shape[0], 3), device="cuda", generator=g)  # 8 seeded per-frame displacement fields, cycled
    torch.cuda.synchronize()
    times = {}
    with torch.cuda.stream(ext):
        for mode in ("rebuild", "refit_only", "rebuild_bound"):
            evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(frames + 5)]
            if mode == "rebuild_bound":
                cs.bind_trs_device(d_pos, d_rot, d_scl)  # zero-copy: b2col_update reads the animated tensors in place
            for f in range(frames + 5):
                evs[f][0].record(ext)
                d_pos.add_(delta[f % 8])  # the frame's animation: every collider moves
                if mode != "rebuild_bound":
                    cs.set_trs_device(d_pos, d_rot, d_scl)
                evs[f][1].record(ext)
                (cs.on_update_refit if mode == "refit_only" else cs.on_update)(wait=False)
                evs[f][2].record(ext)
            cs.sync()
            times[mode] = np.array([[e[1].elapsed_time(e[2]), e[0].elapsed_time(e[2])] for e in evs[5:]])  # [update alone, whole frame]
    st_timed = cs.stats()  # key width / radix passes of the builds that were timed (the policy adapts them to the scene after the first build)
    cs.set_scene(scene)
    cs.on_update()  # leave the tree of the unperturbed scene for the traversal
    ms = float(np.median(times["rebuild"][:, 0]))
    return {"ms": ms, "p99_ms": float(np.percentile(times["rebuild"][:, 0], 99)), "frames": frames,
            "frame_ms": float(np.median(times["rebuild"][:, 1])), "frame_p99_ms": float(np.percentile(times["rebuild"][:, 1], 99)),
            "static_ms": static_ms, "colliders": N, "key_bits": st_timed["key_bits"], "sort_passes": st_timed["sort_passes"],
            "launches_per_update": launches_per_update, "algorithmic_bytes": algo, "achieved_gbs": algo / (ms * 1e-3) / 1e9, "frac": algo / (ms * 1e-3) / 1e9 / peak,
            "static_frac": algo / (static_ms * 1e-3) / 1e9 / peak, "refit_only_ms": float(np.median(times["refit_only"][:, 0])),
            "refit_only_frame_ms": float(np.median(times["refit_only"][:, 1])),
            "bound_trs": {"ms": float(np.median(times["rebuild_bound"][:, 0])), "frame_ms": float(np.median(times["rebuild_bound"][:, 1])),
                          "what": "the same moving scene with the caller's TRS tensors bound (b2col_bind_trs_device): b2col_update reads them in place, the 44 B/entity "
                                  "hand-over copy disappears from the frame (frame_ms = animation + update)"},
            "what": f"every collider moves every frame: a device op perturbs all positions and the TRS arrays are handed over (b2col_set_trs_device, 44 B/entity device to "
                    f"device); ms = b2col_update alone on that moving scene (fused propagate+refit, Morton [{st_timed['key_bits']}-bit keys, chosen by the key-width policy from the collisions the previous build counted], sort [{st_timed['sort_passes']} digits], hierarchy/bounds as one "
                    f"CUDA-graph replay), frame_ms = animation + hand-over + update; medians of {frames} frames, each bracketed by CUDA events on the context's stream; "
                    "static_ms = the same TRS replayed back to back without events in between (round-1 definition)"}


def run_ours(args):
    # This program's stdout is ONE JSON line, but NCCL writes its version banner to the process's stdout when the first
    # communicator comes up: park fd 1 on stderr until the line is printed.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist

    from engine_b200 import build as b2build
    from engine_b200.collision import CollisionSystem, shard_bounds

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — engine_b200 has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa = None
    if world > 1:
        from engine_b200 import dist as b2dist_early

        numa = b2dist_early.bind_to_gpu_numa_node(local)  # before any pinned allocation
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))  # barriers / max-over-ranks / the unique id only
    if rank == 0:
        b2build.build()
    if world > 1:
        dist.barrier()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    cs = CollisionSystem(device=local, use_cuda_graph=True)  # b2col_update replays its kernel sequence as one CUDA graph
    scene = make_scene(args.colliders, args.workload)
    N = scene.n_colliders
    R = args.rays

    # ---- build on rank 0 -------------------------------------------------------------------------------
    rebuild = measure_rebuild(cs, scene, args, torch) if rank == 0 else None

    # ---- replicate over NVLink: NCCL called from C++ inside libb2col.so (b2col_comm_*) --------------------
    multi = None
    if world > 1:
        ids = [CollisionSystem.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        cs.comm_init_rank(world, rank, ids[0])
        barrier()
        t0 = time.perf_counter()
        cs.comm_replicate(0, N)  # cold: first collective of the communicator (connection set-up) + allocation of the replicas' blobs
        cs.sync()
        barrier()
        cold_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)

        def timed_collective(fn, reps=10):
            for _ in range(2):
                fn()
            cs.sync()
            barrier()
            cs.event_record(6)
            for _ in range(reps):
                fn()
            cs.event_record(7)
            cs.sync()
            ms = max_over_ranks(cs.event_elapsed_ms(6, 7)) / reps
            barrier()
            return ms

        blob_bytes = cs.bvh_blob()[1]
        bcast_ms = timed_collective(lambda: cs.comm_replicate(0, N))
        records_ms = timed_collective(lambda: cs.comm_replicate_records(0, N))  # 32 B/collider + Morton/sort/hierarchy on every replica
        cs.comm_replicate(0, N)
        cs.sync()
        barrier()
        multi = {"bvh_blob_bytes": blob_bytes, "bvh_broadcast_ms": bcast_ms, "bvh_broadcast_gbs": blob_bytes / (bcast_ms * 1e-3) / 1e9, "bvh_broadcast_cold_ms": cold_ms,
                 "records_broadcast_plus_rebuild_ms": records_ms, "records_bytes": 32 * N,
                 "how": "warm, 10 back-to-back repeats, CUDA events on every rank's stream, max over ranks: ncclBroadcast straight out of / into the BVH allocations "
                        "(b2col_comm_replicate) vs ncclBroadcast of the 32-byte world records + Morton/sort/hierarchy on every replica (b2col_comm_replicate_records)"}

    # ---- this rank's ray shard (weak scaling: R rays per GPU, the batch is the concatenation in rank order) ----------
    rays_np = make_rays(R, rank, args.workload)
    h_rays = torch.from_numpy(rays_np).pin_memory()
    h_out = torch.empty(R * 36, dtype=torch.uint8).pin_memory()
    d_rays = h_rays.cuda()
    d_out = torch.empty(R * 36, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    # ---- value: device-resident rays and results ----------------------------------------------------
    for _ in range(args.warmup):
        cs.raycast_device(d_rays, R, d_out)
    cs.sync()
    sampler = ClockSampler(local)
    launches0 = cs.stats()["kernel_launches"]
    barrier()
    if rank == 0:
        sampler.start()
    t_wall0 = time.perf_counter()
    cs.event_record(2)
    for _ in range(args.steps):
        cs.raycast_device(d_rays, R, d_out)
    cs.event_record(3)
    cs.sync()
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    launches = cs.stats()["kernel_launches"] - launches0  # kernels launched inside the timed region
    if rank == 0:
        # nvidia-smi needs ~100 ms per sample and the timed region may be shorter: keep the SAME load running (untimed)
        # until at least three samples have been taken under it
        t_wait = time.perf_counter()
        while len(sampler.rows) < 3 and time.perf_counter() - t_wait < 3.0 and sampler.proc is not None:
            cs.raycast_device(d_rays, R, d_out)
            cs.sync()
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = cs.event_elapsed_ms(2, 3)
    step_ms = max_over_ranks(dev_ms) / args.steps
    value = world * R / (step_ms * 1e-3) / 1e6
    hits = int((d_out.view(R, 36)[:, 32] != 0).sum().item())

    # ---- multi-GPU: the same step with the hits gathered on rank 0 INSIDE the timed region ----------------------------
    if world > 1:
        def gather_run(d_shard, n_total, steps):
            for _ in range(2):
                cs.comm_raycast_gather(d_shard, n_total, 0)
            cs.sync()
            barrier()
            cs.event_record(8)
            for _ in range(steps):
                addr = cs.comm_raycast_gather(d_shard, n_total, 0)
            cs.event_record(9)
            cs.sync()
            ms = max_over_ranks(cs.event_elapsed_ms(8, 9)) / steps
            barrier()
            return ms, addr

        # weak: world * R rays, slice r = rank r's rays
        g_ms, addr = gather_run(d_rays, world * R, args.steps)
        multi["value_with_gather"] = world * R / (g_ms * 1e-3) / 1e6
        multi["ms_per_step_with_gather"] = g_ms
        multi["gather_bytes_per_step"] = (world - 1) * R * 36
        # every rank's FULL slice, bitwise, against a 1-GPU trace of the same rays (rank 0 regenerates each rank's rays)
        verified = None
        if rank == 0:
            win = torch.as_tensor(DevMem(addr, world * R * 36), device=f"cuda:{local}")
            verified = True
            tmp_out = torch.empty(R * 36, dtype=torch.uint8, device="cuda")
            for r in range(world):
                d_r = d_rays if r == 0 else torch.from_numpy(make_rays(R, r, args.workload)).cuda()
                cs.raycast_device(d_r, R, tmp_out)
                cs.sync()
                verified = verified and bool(torch.equal(tmp_out, win[r * R * 36 : (r + 1) * R * 36]))
                del d_r
            del tmp_out, win
        multi["all_ranks_bitwise_equal_to_1gpu_trace"] = verified
        # strong: BASELINE configs[2] as written — 16M rays in total, R/N per rank, gathered on rank 0
        S = 16 * 1024 * 1024
        batch = make_rays(S, 0, args.workload)  # the same batch on every rank; a rank uploads its slice only
        lo, hi = shard_bounds(S, world, rank)
        d_slice = torch.from_numpy(batch[lo:hi]).cuda()
        s_ms, addr = gather_run(d_slice, S, max(args.steps, 10))
        strong = {"rays_total": S, "rays_per_gpu": S // world, "ms_per_step_with_gather": s_ms, "value_with_gather": S / (s_ms * 1e-3) / 1e6}
        for _ in range(2):
            cs.raycast_device(d_slice, hi - lo, d_out)
        cs.sync()
        barrier()
        cs.event_record(8)
        for _ in range(max(args.steps, 10)):
            cs.raycast_device(d_slice, hi - lo, d_out)
        cs.event_record(9)
        cs.sync()
        ms_nog = max_over_ranks(cs.event_elapsed_ms(8, 9)) / max(args.steps, 10)
        barrier()
        strong["ms_per_step"] = ms_nog
        strong["value"] = S / (ms_nog * 1e-3) / 1e6
        if rank == 0:
            cs.comm_raycast_gather(d_slice, S, 0)
        else:
            cs.comm_raycast_gather(d_slice, S, 0)
        cs.sync()
        barrier()
        if rank == 0:
            win = torch.as_tensor(DevMem(addr, S * 36), device=f"cuda:{local}")
            d_all = torch.from_numpy(batch).cuda()
            one = torch.empty(S * 36, dtype=torch.uint8, device="cuda")
            cs.raycast_device(d_all, S, one)
            cs.sync()
            strong["bitwise_equal_to_1gpu_trace"] = bool(torch.equal(one, win))
            del win, d_all, one
        multi["strong"] = strong
        multi["comm"] = cs.comm_info()
        del d_slice, batch
        barrier()

    # ---- e2e: the public call with HOST buffers (pinned): H2D + trace + D2H inside the timed region ---
    for _ in range(min(args.warmup, 2)):
        cs.raycast_host_pinned(h_rays, R, h_out)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cs.raycast_host_pinned(h_rays, R, h_out)  # returns when h_out is complete
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / args.steps
    barrier()
    e2e_value = world * R / (e2e_ms * 1e-3) / 1e6
    cs.raycast_device(d_rays, R, d_out)
    cs.sync()
    same = bool(torch.equal(h_out, d_out.cpu()))  # the whole batch, byte for byte

    # ---- e2e, compact results: 12-byte b2col_hit records leave the device instead of 36-byte Raycasts;
    # b2col_expand_hits recomputes location / normal on the host, bit-identical (tests/test_gpu_parity.py), for callers that want them
    compact = None
    if True:
        from engine_b200.capi import ptr as _ptr

        h_hits = torch.empty(R * 12, dtype=torch.uint8).pin_memory()
        for _ in range(2):
            cs.raycast_hits_host_pinned(h_rays, R, h_hits)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            cs.raycast_hits_host_pinned(h_rays, R, h_hits)
        c_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / args.steps
        threads = max(1, min(32, len(os.sched_getaffinity(0)) // world))
        h_out2 = torch.empty(R * 36, dtype=torch.uint8).pin_memory()
        cs.lib.b2col_expand_hits(_ptr(h_rays), _ptr(h_hits), R, _ptr(h_out2), threads)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            cs.raycast_hits_host_pinned(h_rays, R, h_hits)
            cs.lib.b2col_expand_hits(_ptr(h_rays), _ptr(h_hits), R, _ptr(h_out2), threads)
        x_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / args.steps
        barrier()
        compact = {"value": world * R / (c_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": c_ms, "h2d_bytes_per_step": R * 24, "d2h_bytes_per_step": R * 12,
                   "call": "b2col_raycast_hits_batch (entity, distance, hit | side per ray)",
                   "with_host_expansion": {"value": world * R / (x_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": x_ms, "threads_per_rank": threads,
                                           "call": "b2col_raycast_hits_batch + b2col_expand_hits -> the 36-byte Raycast array",
                                           "equal_to_full_format": bool(torch.equal(h_out2, h_out))}}
        del h_hits, h_out2

    if rank != 0:
        if world > 1:
            cs.comm_finalize()
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak_gbs()
    algo_bytes = 60.0 * R + 64.0 * (N - 1)  # 24 B ray in + 36 B Raycast out, node set once
    achieved = algo_bytes / (step_ms * 1e-3) / 1e9
    # dram__bytes_read + dram__bytes_write of ONE launch of this kernel on this workload, from the committed ncu capture of THIS
    # build (tools/capture_traffic.sh writes the file with the hash of the CUDA sources); null when the sources have changed since
    traffic, traffic_note = None, "no ncu capture of this build committed"
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        if (tr["workload"], tr["colliders"], tr["rays"]) == (args.workload, N, R):
            if tr.get("source_hash") == source_hash():
                traffic, traffic_note = tr["dram_bytes_read"] + tr["dram_bytes_write"], f"ncu --set full capture of this build ({tr['source_hash']}), profiles/{tr.get('report', 'r2_traffic.json')}"
            else:
                traffic_note = f"the committed capture belongs to build {tr.get('source_hash')}, this is {source_hash()}: dropped"
    except (OSError, KeyError, ValueError):
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_note,
                "algorithmic_bytes": algo_bytes,
                "kernel": "k_trace_lbvh", "peak_source": peak_src,
                "note": "primary figure = compulsory stream bytes only; traversal at 1M AABBs is L2/latency-bound, not HBM-bound (SURVEY.md 8(d))"}

    # What actually binds the traversal: every node visit is one 32-byte gather per lane, and incoherent
    # rays never share a line, so each visit costs one L1TEX wavefront; B200 sustains ~1.0 such gathers per clock per SM
    # (tools/micro/gather_bench.cu: 0.99).  One extra, untimed pass in the counting mode measures the visits of this batch.
    from engine_b200.capi import MODE_LBVH_COUNT

    c0 = cs.trace_counters()
    cs.raycast_device(d_rays, R, d_out, mode=MODE_LBVH_COUNT)
    cs.sync()
    c1 = cs.trace_counters()
    visits = c1["node_visits"] - c0["node_visits"]
    n_sm = torch.cuda.get_device_properties(local).multi_processor_count
    sm_hz = (clocks or {}).get("sm_mhz") or 1965.0
    per_clk_sm = visits / (step_ms * 1e-3) / (n_sm * sm_hz * 1e6)
    roofline["binding"] = {"resource": "L1TEX wavefronts: one 32-byte node gather per lane per visit", "node_visits_per_ray": visits / R,
                           "achieved": per_clk_sm, "peak": 1.0, "unit": "node gathers/clk/SM", "frac": per_clk_sm / 1.0,
                           "l2_to_sm_gbs": visits * 32.0 / (step_ms * 1e-3) / 1e9,
                           "peak_source": "measured gather rate of an L2-resident table, tools/micro/gather_bench.cu (0.99); stack and ray traffic in shared memory share the pipe"}

    extras = {}
    if world == 1 and not args.no_extras:
        del d_rays, d_out, h_rays, h_out
        torch.cuda.empty_cache()
        from tools import bench_extras

        extras = bench_extras.run_all(cs, args, torch)

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        arm = CpuArm(scene)
        n = arm.calibrate(rays_np, target_s=12.0)
        secs = arm.trace(rays_np[:n])
        cpu_baseline = {"value": n / secs / 1e6, "unit": UNIT, "cores": arm.threads, "kind": arm.kind,
                        "sample": f"first {n} rays of the same batch, one pass, {arm.kind} GetRaycast on a binned-SAH tree injected into bvh_ (reference BuildNode is O(N^2): 1M not buildable)",
                        "setup_s": arm.setup_s}
        # the reference is single-threaded (src/application.cpp:390): the faithful figure, on a smaller sample
        n1 = max(20000, min(n // (4 * arm.threads), n))
        threads, arm.threads = arm.threads, 1
        secs1 = arm.trace(rays_np[:n1])
        arm.threads = threads
        cpu_baseline["single_thread"] = {"value": n1 / secs1 / 1e6, "unit": UNIT, "sample": f"first {n1} rays of the same batch, one thread"}

    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": R * 24, "d2h_bytes_per_step": R * 36, "ms_per_step": e2e_ms,
                "matches_device_path": same, "checked": "every byte of the batch", "compact_results": compact},
        "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu_baseline, "clocks": clocks,
        "rebuild": rebuild, "numa_cpulist_rank0": numa, "hit_fraction": hits / R, "wall_ms_timed_region": wall_ms,
    }
    if multi:
        line["multi_gpu"] = multi
        line["value_with_gather"] = multi["value_with_gather"]
        line["bvh_broadcast_ms"] = multi["bvh_broadcast_ms"]
    line.update(extras)
    print(json.dumps(line))
    if world > 1:
        cs.comm_finalize()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--colliders", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=16 * 1024 * 1024, help="rays per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra 1-GPU records (config 4 dynamic frames, config 5 aimed rays, update through the drop-in)")
    ap.add_argument("--workload", default="config3", choices=["config3", "config5"], help="config3: uniform 1M (the headline); config5: clustered (use --colliders 10000000)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
