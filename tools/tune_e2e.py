"""dev tool: sweep the chunk schedule of the host-buffer pipeline (B2COL_E2E_FIRST_CHUNK / B2COL_E2E_CHUNK)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = r'''
import sys, time, torch
sys.path.insert(0, %r)
from engine_b200 import synth
from engine_b200.collision import CollisionSystem
cs = CollisionSystem(use_cuda_graph=True)
cs.set_scene(synth.uniform_boxes(1_000_000)); cs.on_update()
for R in (1 << 20, 16 << 20):
    rays = synth.random_rays(R, seed=4)
    h_rays = torch.from_numpy(rays).pin_memory(); h_out = torch.empty(R * 36, dtype=torch.uint8).pin_memory()
    for _ in range(2): cs.raycast_host_pinned(h_rays, R, h_out)
    t0 = time.perf_counter()
    K = 5
    for _ in range(K): cs.raycast_host_pinned(h_rays, R, h_out)
    ms = (time.perf_counter() - t0) * 1e3 / K
    print(f"  R={R:9d}  {ms:7.3f} ms  {R / ms / 1e3:7.1f} Mrays/s", flush=True)
''' % ROOT
for first, big in [(128, 2048), (256, 1024), (128, 512), (256, 512), (512, 1024), (256, 256), (512, 512), (64, 256)]:
    env = dict(os.environ, B2COL_E2E_FIRST_CHUNK=str(first << 10), B2COL_E2E_CHUNK=str(big << 10))
    print(f"first {first}K max {big}K", flush=True)
    subprocess.run([sys.executable, "-c", code], env=env)
