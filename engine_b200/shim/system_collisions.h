// Drop-in replacement for the reference's include/system_collisions.h (bailwillharr/engine).
//
// Same namespace, class name, base class, public method names and public `bvh_` member as the
// reference class (reference include/system_collisions.h:28-74), so that scene.cpp (:29 registers
// it), application.cpp (:61-62,153-154 read bvh_) and game code calling
// scene->GetSystem<CollisionSystem>()->GetRaycast(ray) compile unchanged when this directory is put
// in front of the reference's include path.  The bodies (system_collisions.cpp next to this file)
// forward to the C ABI of libb2col.so (include/b2col.h): the BVH lives on the GPU.
//
// Additive API: GetRaycasts (batch), GetRaycastHits (batch, compact results), UseDevices (several GPUs),
// SyncDebugBvh / keep_debug_bvh (materialise bvh_ on demand).
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

#include <glm/vec3.hpp>

#include "component_collider.h"
#include "ecs.h"

struct b2col_ctx;  // include/b2col.h

namespace engine {

// Layouts are part of the contract with libb2col.so: static_asserts in system_collisions.cpp.
struct Ray {
    glm::vec3 origin;
    glm::vec3 direction;  // need not be normalised
};

struct Raycast {
    glm::vec3 location;
    glm::vec3 normal;
    Entity hit_entity;
    float distance;  // negative when the origin is inside the hit box, as in the reference
    bool hit;
};

enum class AABBSide { Left, Right, Bottom, Top, Front, Back };

class CollisionSystem : public System {
   public:
    explicit CollisionSystem(Scene* scene);
    ~CollisionSystem() override;

    void onComponentInsert(Entity entity) override;

    // Gathers ColliderComponent::aabb + TransformComponent::world_matrix from the ECS, uploads them and
    // enqueues refit + LBVH build on the GPU (the reference rebuilds every frame too).
    void onUpdate(float ts) override;

    // Nearest hit, reference semantics. One ray = a batch of one (latency-bound; prefer GetRaycasts).
    Raycast GetRaycast(Ray ray);
    void GetRaycasts(const Ray* rays, std::size_t count, Raycast* results);
    // Compact results: 12 bytes per ray cross PCIe instead of 36 (hit entity, distance, hit flag and box side); ExpandHits
    // recomputes location and normal on the host, bit-identical to GetRaycasts (threads = host threads to use).
    struct RaycastHit {
        Entity hit_entity;
        float distance;
        uint32_t hit_and_side;  // bit 0: hit; bits 8-10: AABBSide
    };
    void GetRaycastHits(const Ray* rays, std::size_t count, RaycastHit* results);
    static void ExpandHits(const Ray* rays, const RaycastHit* hits, std::size_t count, Raycast* results, int threads = 1);

    // Several GPUs of this machine (also: environment B2COL_DEVICES=all|<count>|<id>,<id>,...): the first device builds, the
    // tree is replicated with one NCCL broadcast after an update that a large batch follows, and GetRaycasts batches of at
    // least multi_gpu_min_rays are cut into one contiguous slice per GPU (results bit-identical to one GPU).
    void UseDevices(const std::vector<int>& devices);
    std::size_t multi_gpu_min_rays = 262144;
    int devices_in_use() const;

    // One node of the BVH in the reference's exchange format (60 bytes, align 4).
    struct BiTreeNode {
        enum class Type : uint8_t { BoundingVolume, Entity, Empty };
        AABB box1;
        AABB box2;
        uint32_t index1;
        uint32_t index2;
        Type type1;
        Type type2;
    };

    // The debug overlay reads this member directly. It is refreshed from the GPU tree by
    // SyncDebugBvh(), or after every onUpdate when keep_debug_bvh is set (costs a device->host copy).
    std::vector<BiTreeNode> bvh_{};
    bool keep_debug_bvh = false;
    void SyncDebugBvh();

    // Full rebuild every `rebuild_interval` frames; the frames in between only refit (b2col_refit: same tree
    // shape, new boxes — the reference's own TODO, src/system_collisions.cpp:116-120).  1 = the reference's
    // behaviour, rebuild every frame (:121).  A frame whose collider set (entities or local AABBs) differs from
    // the previous frame's always rebuilds.  Query results do not depend on this setting except for the winner
    // among exactly tied hits.
    unsigned rebuild_interval = 1;
    // ColliderComponent::aabb is compared with the uploaded copy every collider_check_interval-th frame (1 = every frame, the
    // reference's semantics: it re-reads every box every frame; 0 = never: boxes are immutable once added).  New colliders are
    // always picked up at once (onComponentInsert).
    unsigned collider_check_interval = 1;

   private:
    struct B2colSceneCtx* shared_ = nullptr;
    b2col_ctx* ctx_ = nullptr;
    std::size_t colliders_inserted_ = 0;
    std::size_t seen_inserts_ = 0;
    std::vector<float> gather_world_;
    std::vector<uint32_t> device_entities_;  // the collider set currently on the device, ascending entity IDs
    std::vector<AABB> device_boxes_;         // ... and its local boxes, same order
    unsigned frames_since_check_ = 0;
    bool BoxesUnchanged(const ColliderComponent* base) const;
    unsigned frames_since_build_ = 0;
    bool have_build_ = false;
    void Check(int status, const char* what) const;
};

}  // namespace engine
