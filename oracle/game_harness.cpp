// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// Runs the test game's player controller headless inside the reference's own Scene.
// The controller is the UNMODIFIED reference test/src/camera_controller.cpp, compiled where it lies (oracle/Makefile)
// against the scripted-input stand-ins in stubs/game/ for application.h / input_manager.h / window.h /
// scene_manager.h.  Built three ways, like ref_harness.cpp: with the reference's CollisionSystem (the oracle),
// and with engine_b200/shim's drop-in classes, where the batched port
// (engine_b200/shim/game/camera_controller_batched.cpp) can be registered instead of the original.
#include <cstring>

#include "application.h"  // stubs/game
#include "camera_controller.hpp"
#include "harness_types.h"
#include "input_manager.h"
#include "system_collisions.h"
#ifdef B2COL_DROPIN
#include "camera_controller_batched.hpp"
#endif

using namespace engine;

extern "C" {

// Mirrors reference test/src/game.cpp:92-109: a "camera" entity at `pos3`, the controller component, and the
// controller system registered at index 0 — it runs BEFORE Transform/Collision, so a frame's queries see the
// tree of the previous frame (frame 0: empty tree, everything misses).
// batched != 0 registers the batched port instead (drop-in builds only; returns 0 where unavailable).
uint32_t gameh_add_camera(void* h, const float* pos3, int batched)
{
    auto* H = static_cast<Harness*>(h);
    const Entity camera = H->scene.CreateEntity("camera");
    H->scene.GetComponent<TransformComponent>(camera)->position = {pos3[0], pos3[1], pos3[2]};
    H->scene.RegisterComponent<CameraControllerComponent>();
    if (batched) {
#ifdef B2COL_DROPIN
        H->scene.RegisterSystemAtIndex<BatchedCameraControllerSystem>(0);
#else
        return 0;
#endif
    }
    else {
        H->scene.RegisterSystemAtIndex<CameraControllerSystem>(0);
    }
    H->scene.AddComponent<CameraControllerComponent>(camera);
    H->camera = camera;
    return camera;
}

// n_frames of Scene::Update(dt).  input: 6 floats per frame = movex, movey, lookx, looky, sprint, jump.
// state out: 16 floats per frame = position xyz, velocity xyz, rotation xyzw, yaw, pitch, grounded, 3 spare (0).
void gameh_run(void* h, size_t n_frames, const float* input6, float dt, float* state16)
{
    auto* H = static_cast<Harness*>(h);
    InputManager* in = H->app->getInputManager();
    auto* t = H->scene.GetComponent<TransformComponent>(H->camera);
    auto* c = H->scene.GetComponent<CameraControllerComponent>(H->camera);
    for (size_t f = 0; f < n_frames; ++f) {
        const float* i6 = input6 + 6 * f;
        in->movex = i6[0];
        in->movey = i6[1];
        in->lookx = i6[2];
        in->looky = i6[3];
        in->sprint = i6[4] != 0.0f;
        in->jump = i6[5] != 0.0f;
        H->scene.Update(dt);
        H->app->debug_lines.clear();  // the renderer would consume these every frame
        t = H->scene.GetComponent<TransformComponent>(H->camera);
        float* s = state16 + 16 * f;
        std::memset(s, 0, 64);
        s[0] = t->position.x; s[1] = t->position.y; s[2] = t->position.z;
        s[3] = c->vel.x; s[4] = c->vel.y; s[5] = c->vel.z;
        s[6] = t->rotation.x; s[7] = t->rotation.y; s[8] = t->rotation.z; s[9] = t->rotation.w;
        s[10] = c->yaw; s[11] = c->pitch; s[12] = c->grounded ? 1.0f : 0.0f;
    }
}

// GetRaycasts calls, rays, frames, re-casts of the batched controller (zeros otherwise)
void gameh_batched_counters(void* h, uint64_t* out4)
{
    out4[0] = out4[1] = out4[2] = out4[3] = 0;
#ifdef B2COL_DROPIN
    auto* H = static_cast<Harness*>(h);
    try {
        auto* s = H->scene.GetSystem<BatchedCameraControllerSystem>();
        out4[0] = s->raycast_calls;
        out4[1] = s->rays_cast;
        out4[2] = s->frames;
        out4[3] = s->recasts;
    }
    catch (...) {
    }
#endif
}

#ifdef B2COL_DROPIN
void gameh_set_rebuild_interval(void* h, unsigned interval)
{
    static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>()->rebuild_interval = interval;
}
#endif

}  // extern "C"
