// Batched port of the test game's player controller.
//
// The reference's CameraControllerSystem::onUpdate (reference test/src/camera_controller.cpp:26-240) issues its
// collision queries one GetRaycast at a time: kNumHorizontalRays = 20 coplanar wall probes (:77-97), then one
// vertical probe, down while falling (:119-124) or up while jumping (:143-149).  Against the GPU backend each
// single-ray call is a full launch + two copies, so this system asks for the frame's 21 rays in ONE
// CollisionSystem::GetRaycasts call.  The vertical probe starts from the player's position AFTER the wall
// response, so it is cast speculatively from the unpushed position and re-cast alone on the (rare) frames where
// a wall pushed the player — the tree does not change inside a frame, so the answers equal the sequential ones.
//
// It reuses the game's own CameraControllerComponent (reference test/src/camera_controller.hpp:11-42) and the
// engine's input interface, and is registered exactly like the original (reference test/src/game.cpp:107-109):
//     scene->RegisterSystemAtIndex<BatchedCameraControllerSystem>(0);
// Window-key handling (:186-237: reset, noclip, gravity toggle, click-to-pick) is UI, not collision response, and
// stays with the original system.
#pragma once

#include <cstddef>

#include "camera_controller.hpp"  // the game's component + constants
#include "system_collisions.h"    // engine_b200/shim: GetRaycasts

class BatchedCameraControllerSystem : public engine::System {
   public:
    explicit BatchedCameraControllerSystem(engine::Scene* scene);
    void onUpdate(float ts) override;

    // GetRaycasts calls / rays issued so far (tests: at most 2 calls per frame, 1 on frames without a wall push)
    std::size_t raycast_calls = 0, rays_cast = 0, frames = 0, recasts = 0;

   private:
    engine::TransformComponent* transform_ = nullptr;
    CameraControllerComponent* state_ = nullptr;
    engine::CollisionSystem* collisions_ = nullptr;
    bool Bind();
};
