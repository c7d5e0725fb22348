// One b2col context per engine::Scene, shared by the drop-in TransformSystem and CollisionSystem so that
// world matrices computed on the GPU by the former are refitted by the latter without leaving the device.
#pragma once
#include <vector>

#include "b2col.h"

namespace engine {
class Scene;

struct B2colSceneCtx {
    b2col_ctx* ctx = nullptr;
    int users = 0;
    bool transforms_on_device = false;  // set by the drop-in TransformSystem after b2col_propagate
    unsigned n_entities = 0;            // entity rows currently on the device
    // Scale-out of GetRaycasts over the GPUs of this process (north_star: "sharded ray batches"): ranks[0] == ctx builds the
    // tree, the others receive it with ONE NCCL broadcast after every update that a large batch follows (lazily:
    // frames without a large batch never pay for it).  Empty = single GPU.
    std::vector<b2col_ctx*> ranks;
    bool replicas_fresh = false;  // the other ranks hold the tree of the last update
};

// B2COL_DEVICES=all | <count> | <id>,<id>,...  (read at the first Acquire; the first device builds).  Also callable directly.
// Throws std::runtime_error when a device or NCCL is unavailable.
void B2colUseDevices(B2colSceneCtx* shared, const int* devices, int n);

// Creates the context on first use (throws std::runtime_error when there is no CUDA device: no CPU path).
B2colSceneCtx* B2colAcquire(Scene* scene);
void B2colRelease(Scene* scene);
}  // namespace engine
