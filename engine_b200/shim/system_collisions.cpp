// Bodies of the drop-in engine::CollisionSystem: ECS gather + calls into libb2col.so (include/b2col.h).
// Compile this file INSTEAD of the reference's src/system_collisions.cpp and link libb2col.so.
#include "system_collisions.h"

#include <cstring>
#include <stdexcept>
#include <string>
#include <typeinfo>

#include "b2col.h"
#include "b2col_scene_ctx.h"
#include "component_transform.h"
#include "scene.h"

namespace engine {

static_assert(sizeof(Ray) == sizeof(b2col_ray) && sizeof(Raycast) == sizeof(b2col_raycast), "query PODs must match the C ABI");
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

void CollisionSystem::onUpdate(float ts)
{
    (void)ts;
    // world matrices of every entity, indexed by entity ID (row 0 unused) — what TransformSystem just wrote
    const uint32_t n_entities = m_scene->next_entity_id_ - 1u;
    gather_entities_.clear();
    gather_boxes_.clear();
    for (Entity entity : m_entities) {  // ascending IDs (std::set), the reference's own iteration order
        const auto* c = m_scene->GetComponent<ColliderComponent>(entity);
        gather_entities_.push_back(entity);
        gather_boxes_.push_back(c->aabb);
    }
    if (shared_->transforms_on_device && shared_->n_entities == n_entities) {
        // the drop-in TransformSystem of this Scene has already put this frame's world matrices on the device
    }
    else {
        // reference TransformSystem on the CPU: upload TransformComponent::world_matrix.  Only collider entities'
        // matrices are read by the refit, so only those rows are filled.
        gather_world_.resize(16u * (static_cast<std::size_t>(n_entities) + 1u));
        for (Entity entity : gather_entities_) {
            const auto* t = m_scene->GetComponent<TransformComponent>(entity);
            std::memcpy(&gather_world_[16u * entity], &t->world_matrix, 64);
        }
        Check(b2col_set_world_matrices(ctx_, n_entities, gather_world_.data()), "b2col_set_world_matrices");
    }
    const bool same_set = have_build_ && gather_entities_ == device_entities_ &&
                          (gather_boxes_.empty() || std::memcmp(gather_boxes_.data(), device_boxes_.data(), gather_boxes_.size() * sizeof(AABB)) == 0);
    if (same_set && rebuild_interval > 1 && frames_since_build_ + 1 < rebuild_interval) {
        Check(b2col_refit(ctx_), "b2col_refit");
        ++frames_since_build_;
    }
    else {
        if (!same_set) {
            Check(b2col_set_colliders(ctx_, static_cast<uint32_t>(gather_entities_.size()), gather_entities_.data(),
                                      reinterpret_cast<const b2col_aabb*>(gather_boxes_.data())),
                  "b2col_set_colliders");
            device_entities_ = gather_entities_;
            device_boxes_ = gather_boxes_;
        }
        Check(b2col_update(ctx_), "b2col_update");
        have_build_ = true;
        frames_since_build_ = 0;
    }
    if (keep_debug_bvh) SyncDebugBvh();
}

void CollisionSystem::GetRaycasts(const Ray* rays, std::size_t count, Raycast* results)
{
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
    bvh_.resize(2u * gather_entities_.size() + 1u);
    std::size_t n = 0;
    Check(b2col_export_bitreenode(ctx_, reinterpret_cast<b2col_bitreenode*>(bvh_.data()), bvh_.size(), &n), "b2col_export_bitreenode");
    bvh_.resize(n);
}

}  // namespace engine
