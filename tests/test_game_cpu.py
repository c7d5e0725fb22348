"""The test game's player controller, headless: the UNMODIFIED reference
test/src/camera_controller.cpp inside the reference Scene with the reference CollisionSystem (oracle/game_harness.cpp)
reproduces the committed golden trajectory.  (The GPU-side comparison — drop-in CollisionSystem and the batched
port — is in test_gpu_parity.py.)"""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from engine_b200 import synth
from oracle.oracle import RefEngine


@pytest.mark.skipif(not RefEngine.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_reference_game_reproduces_golden_trajectory():
    z = np.load(os.path.join(GOLDEN, "game_traj.npz"))
    assert (z["input"] == synth.game_inputs(600)).all(), "input generator drifted from the fixture"
    s = synth.game_scene(140)
    e = RefEngine()
    e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
    e.add_colliders(s.collider_entity, s.collider_aabb)
    e.add_camera(z["camera"])
    state = e.run_game(z["input"][:300])
    assert (state.view(np.uint32) == z["state"][:300].view(np.uint32)).all()
    assert state[:, 12].sum() > 50  # the player spends time standing on colliders
