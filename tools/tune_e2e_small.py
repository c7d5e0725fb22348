"""dev tool: e2e (pinned host buffers) throughput of mid-size batches for several B2COL_E2E_MIN_CHUNK settings
(256K = the schedule before the adaptive first chunk: pipelining only from 512K rays up)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = r'''
import sys, time, json, os, torch
sys.path.insert(0, %r)
from engine_b200 import synth
from engine_b200.collision import CollisionSystem
cs = CollisionSystem(use_cuda_graph=True)
cs.set_scene(synth.uniform_boxes(1_000_000)); cs.on_update()
rays_all = synth.random_rays(4 << 20, seed=4)
out = {}
for R in (32 << 10, 64 << 10, 128 << 10, 256 << 10, 512 << 10, 1 << 20, 2 << 20, 4 << 20):
    h_rays = torch.from_numpy(rays_all[:R].copy()).pin_memory(); h_out = torch.empty(R * 36, dtype=torch.uint8).pin_memory()
    for _ in range(3): cs.raycast_host_pinned(h_rays, R, h_out)
    K = 20
    t0 = time.perf_counter()
    for _ in range(K): cs.raycast_host_pinned(h_rays, R, h_out)
    ms = (time.perf_counter() - t0) * 1e3 / K
    out[R] = round(R / ms / 1e3, 1)
print(json.dumps({"min_chunk": int(os.environ["B2COL_E2E_MIN_CHUNK"]), "e2e_Mrays_s_by_batch": out}), flush=True)
''' % ROOT
for mc in (256, 64, 32, 16, 8):
    env = dict(os.environ, B2COL_E2E_MIN_CHUNK=str(mc << 10))
    subprocess.run([sys.executable, "-c", code], env=env)
