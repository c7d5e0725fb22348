// Bodies of the drop-in engine::CollisionSystem: ECS gather + calls into libb2col.so (include/b2col.h).
// Compile this file INSTEAD of the reference's src/system_collisions.cpp and link libb2col.so.
#include "system_collisions.h"

#include <cstring>
#include <stdexcept>
#include <string>
#include <typeinfo>

#include <atomic>

#include "b2col.h"
#include "b2col_scene_ctx.h"
#include "b2col_workers.h"
#include "component_transform.h"
#include "scene.h"

namespace engine {

static_assert(sizeof(Ray) == sizeof(b2col_ray) && sizeof(Raycast) == sizeof(b2col_raycast), "query PODs must match the C ABI");
static_assert(sizeof(CollisionSystem::RaycastHit) == sizeof(b2col_hit), "compact results must match the C ABI");
static_assert(sizeof(AABB) == sizeof(b2col_aabb) && sizeof(CollisionSystem::BiTreeNode) == sizeof(b2col_bitreenode), "tree PODs must match the C ABI");

CollisionSystem::CollisionSystem(Scene* scene)
    : System(scene, {typeid(TransformComponent).hash_code(), typeid(ColliderComponent).hash_code()})
{
    // error convention of the reference: exceptions (caught in main, reference test/src/main.cpp:29-35)
    shared_ = B2colAcquire(scene);  // one context per Scene, shared with the drop-in TransformSystem if that is in use
    ctx_ = shared_->ctx;
}

CollisionSystem::~CollisionSystem() { B2colRelease(m_scene); }

void CollisionSystem::Check(int status, const char* what) const
{
    if (status != B2COL_OK) throw std::runtime_error(std::string(what) + ": " + b2col_last_error(ctx_));
}

void CollisionSystem::onComponentInsert(Entity entity)
{
    (void)entity;
    ++colliders_inserted_;
}

// Did any collider's local AABB change since it was uploaded?  ColliderComponent has no dirty flag and the reference re-reads
// every box every frame (src/system_collisions.cpp:125-127), so the check is exact — but it runs as memcmp over RUNS of
// consecutive entity IDs (the component array is dense in entity ID, include/ecs.h:28-49, and device_boxes_ is in the same
// order), not as a per-entity component lookup: ~2 ms per million colliders instead of ~100.
bool CollisionSystem::BoxesUnchanged(const ColliderComponent* base) const
{
    static_assert(sizeof(ColliderComponent) == sizeof(AABB), "ColliderComponent is its AABB: runs of entities are runs of boxes");
    const std::size_t n = device_entities_.size();
    std::atomic<bool> same{true};
    B2colWorkers::Get().ParallelFor(n, 65536, [&](std::size_t lo, std::size_t hi) {
        std::size_t i = lo;
        while (i < hi && same.load(std::memory_order_relaxed)) {
            std::size_t j = i + 1;
            while (j < hi && device_entities_[j] == device_entities_[j - 1] + 1u) ++j;
            if (std::memcmp(&base[device_entities_[i]], &device_boxes_[i], (j - i) * sizeof(AABB)) != 0) same.store(false, std::memory_order_relaxed);
            i = j;
        }
    });
    return same.load();
}

void CollisionSystem::onUpdate(float ts)
{
    (void)ts;
    // world matrices of every entity, indexed by entity ID (row 0 unused) — what TransformSystem just wrote
    const uint32_t n_entities = m_scene->next_entity_id_ - 1u;
    const ColliderComponent* cbase = nullptr;
    if (!m_entities.empty()) {
        const Entity first = *m_entities.begin();
        cbase = m_scene->GetComponent<ColliderComponent>(first) - first;  // one lookup: the array is a dense vector indexed by entity ID
    }
    // the collider set only ever grows (the reference has no component removal, include/ecs.h:41-44), and only through
    // onComponentInsert: the sorted entity list is rebuilt when that fired, the boxes are compared every
    // collider_check_interval-th frame
    bool same_set = have_build_ && seen_inserts_ == colliders_inserted_ && device_entities_.size() == m_entities.size();
    if (same_set && cbase && collider_check_interval != 0 && ++frames_since_check_ >= collider_check_interval) {
        frames_since_check_ = 0;
        same_set = BoxesUnchanged(cbase);
    }
    if (!same_set) {
        device_entities_.assign(m_entities.begin(), m_entities.end());  // ascending IDs (std::set), the reference's own iteration order
        device_boxes_.resize(device_entities_.size());
        B2colWorkers::Get().ParallelFor(device_entities_.size(), 65536, [&](std::size_t lo, std::size_t hi) {
            for (std::size_t i = lo; i < hi; ++i) device_boxes_[i] = cbase[device_entities_[i]].aabb;
        });
        seen_inserts_ = colliders_inserted_;
        frames_since_check_ = 0;
    }
    if (shared_->transforms_on_device && shared_->n_entities == n_entities) {
        // the drop-in TransformSystem of this Scene has already put this frame's world matrices on the device
    }
    else if (!m_entities.empty()) {
        // reference TransformSystem on the CPU: upload TransformComponent::world_matrix.  Only collider entities'
        // matrices are read by the refit, so only those rows are filled.
        gather_world_.resize(16u * (static_cast<std::size_t>(n_entities) + 1u));
        const Entity first = device_entities_.front();
        const TransformComponent* tbase = m_scene->GetComponent<TransformComponent>(first) - first;
        B2colWorkers::Get().ParallelFor(device_entities_.size(), 16384, [&](std::size_t lo, std::size_t hi) {
            for (std::size_t i = lo; i < hi; ++i) std::memcpy(&gather_world_[16u * device_entities_[i]], &tbase[device_entities_[i]].world_matrix, 64);
        });
        Check(b2col_set_world_matrices(ctx_, n_entities, gather_world_.data()), "b2col_set_world_matrices");
    }
    if (same_set && rebuild_interval > 1 && frames_since_build_ + 1 < rebuild_interval) {
        Check(b2col_refit(ctx_), "b2col_refit");
        ++frames_since_build_;
    }
    else {
        if (!same_set) {
            Check(b2col_set_colliders(ctx_, static_cast<uint32_t>(device_entities_.size()), device_entities_.data(),
                                      reinterpret_cast<const b2col_aabb*>(device_boxes_.data())),
                  "b2col_set_colliders");
        }
        Check(b2col_update(ctx_), "b2col_update");
        have_build_ = true;
        frames_since_build_ = 0;
    }
    shared_->replicas_fresh = false;  // the other GPUs (UseDevices) still hold the previous tree
    if (keep_debug_bvh) SyncDebugBvh();
}

void CollisionSystem::UseDevices(const std::vector<int>& devices)
{
    // devices[0] must be the device this Scene's context was created on (B2COL_DEVICES chooses it; default 0)
    B2colUseDevices(shared_, devices.data(), static_cast<int>(devices.size()));
}

int CollisionSystem::devices_in_use() const { return shared_->ranks.empty() ? 1 : static_cast<int>(shared_->ranks.size()); }

void CollisionSystem::GetRaycastHits(const Ray* rays, std::size_t count, RaycastHit* results)
{
    Check(b2col_raycast_hits_batch(ctx_, reinterpret_cast<const b2col_ray*>(rays), count, reinterpret_cast<b2col_hit*>(results)), "b2col_raycast_hits_batch");
}

void CollisionSystem::ExpandHits(const Ray* rays, const RaycastHit* hits, std::size_t count, Raycast* results, int threads)
{
    b2col_expand_hits(reinterpret_cast<const b2col_ray*>(rays), reinterpret_cast<const b2col_hit*>(hits), count, reinterpret_cast<b2col_raycast*>(results), threads);
}

void CollisionSystem::GetRaycasts(const Ray* rays, std::size_t count, Raycast* results)
{
    if (shared_->ranks.size() > 1 && count >= multi_gpu_min_rays && have_build_) {
        const int n = static_cast<int>(shared_->ranks.size());
        if (!shared_->replicas_fresh) {
            Check(b2col_comm_replicate_all(shared_->ranks.data(), n, 0), "b2col_comm_replicate_all");
            shared_->replicas_fresh = true;
        }
        Check(b2col_mgpu_raycast_batch(shared_->ranks.data(), n, reinterpret_cast<const b2col_ray*>(rays), count, reinterpret_cast<b2col_raycast*>(results), B2COL_MODE_LBVH),
              "b2col_mgpu_raycast_batch");
        return;
    }
    Check(b2col_raycast_batch(ctx_, reinterpret_cast<const b2col_ray*>(rays), count, reinterpret_cast<b2col_raycast*>(results), B2COL_MODE_LBVH),
          "b2col_raycast_batch");
}

Raycast CollisionSystem::GetRaycast(Ray ray)
{
    Raycast res{};
    GetRaycasts(&ray, 1, &res);
    return res;
}

void CollisionSystem::SyncDebugBvh()
{
    bvh_.resize(2u * device_entities_.size() + 1u);
    std::size_t n = 0;
    Check(b2col_export_bitreenode(ctx_, reinterpret_cast<b2col_bitreenode*>(bvh_.data()), bvh_.size(), &n), "b2col_export_bitreenode");
    bvh_.resize(n);
}

}  // namespace engine
