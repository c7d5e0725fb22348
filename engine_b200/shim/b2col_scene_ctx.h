// One b2col context per engine::Scene, shared by the drop-in TransformSystem and CollisionSystem so that
// world matrices computed on the GPU by the former are refitted by the latter without leaving the device.
#pragma once
#include "b2col.h"

namespace engine {
class Scene;

struct B2colSceneCtx {
    b2col_ctx* ctx = nullptr;
    int users = 0;
    bool transforms_on_device = false;  // set by the drop-in TransformSystem after b2col_propagate
    unsigned n_entities = 0;            // entity rows currently on the device
};

// Creates the context on first use (throws std::runtime_error when there is no CUDA device: no CPU path).
B2colSceneCtx* B2colAcquire(Scene* scene);
void B2colRelease(Scene* scene);
}  // namespace engine
