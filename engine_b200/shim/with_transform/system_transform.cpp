// Bodies of the drop-in engine::TransformSystem (replaces reference src/system_transform.cpp).
#include "system_transform.h"

#include <cstring>
#include <stdexcept>
#include <string>
#include <typeinfo>

#include "b2col.h"
#include "b2col_scene_ctx.h"
#include "component_transform.h"
#include "scene.h"

namespace engine {

static_assert(sizeof(glm::quat) == 16 && sizeof(glm::vec3) == 12 && sizeof(glm::mat4) == 64, "glm layouts the packing relies on");

TransformSystem::TransformSystem(Scene* scene) : System(scene, {typeid(TransformComponent).hash_code()}) { shared_ = B2colAcquire(scene); }

TransformSystem::~TransformSystem() { B2colRelease(m_scene); }

void TransformSystem::Check(int status, const char* what) const
{
    if (status != B2COL_OK) throw std::runtime_error(std::string(what) + ": " + b2col_last_error(shared_->ctx));
}

void TransformSystem::onUpdate(float ts)
{
    (void)ts;
    if (m_entities.empty()) return;
    const uint32_t n_entities = m_scene->next_entity_id_ - 1u;
    const std::size_t rows = static_cast<std::size_t>(n_entities) + 1u;
    const bool grown = parent_.size() != rows;
    if (grown) {  // new entities: rows that hold no transform stay identity roots
        parent_.assign(rows, 0u);
        pos_.assign(3u * rows, 0.0f);
        rot_.assign(4u * rows, 0.0f);
        scl_.assign(3u * rows, 1.0f);
        static_.assign(rows, 0);
        world_.resize(16u * rows);
        for (std::size_t e = 0; e < rows; ++e) rot_[4u * e + 3u] = 1.0f;
        packed_once_ = false;
    }
    // The component array is one dense vector indexed by entity ID (reference include/ecs.h:28-49), so one
    // lookup yields the base pointer instead of two hash lookups + a dynamic_cast per entity
    // (include/scene.h:56-69,158-170 — the cost that dominates the reference's TransformSystem).
    const Entity first = *m_entities.begin();
    TransformComponent* const base = m_scene->GetComponent<TransformComponent>(first) - first;

    // ---- gather: AoS -> SoA (static rows only the first time)
    const bool all_rows = !(skip_static_rows && packed_once_);
    for (Entity e : m_entities) {
        const TransformComponent& t = base[e];
        if (!all_rows && static_[e] && t.is_static) continue;
        parent_[e] = t.parent;
        std::memcpy(&pos_[3u * e], &t.position, 12);
        std::memcpy(&rot_[4u * e], &t.rotation, 16);  // glm::quat memory order is x,y,z,w
        std::memcpy(&scl_[3u * e], &t.scale, 12);
        static_[e] = t.is_static ? 1 : 0;
    }
    if (grown || parent_ != parent_prev_) {
        Check(b2col_set_hierarchy(shared_->ctx, n_entities, parent_.data()), "b2col_set_hierarchy");
        parent_prev_ = parent_;
    }
    Check(b2col_set_trs(shared_->ctx, n_entities, pos_.data(), rot_.data(), scl_.data()), "b2col_set_trs");
    Check(b2col_propagate(shared_->ctx), "b2col_propagate");
    shared_->transforms_on_device = true;
    shared_->n_entities = n_entities;

    // ---- scatter: world_matrix back into the components
    Check(b2col_get_world_matrices(shared_->ctx, 0u, n_entities + 1u, world_.data()), "b2col_get_world_matrices");
    for (Entity e : m_entities) {
        TransformComponent& t = base[e];
        if (!all_rows && static_[e]) continue;  // unchanged since the first frame
        std::memcpy(&t.world_matrix, &world_[16u * e], 64);
    }
    packed_once_ = true;
}

Entity TransformSystem::GetChildEntity(Entity parent, const std::string& tag)
{
    for (Entity entity : m_entities) {
        const auto* t = m_scene->GetComponent<TransformComponent>(entity);
        if (t->parent == parent && t->tag == tag) return entity;
    }
    return 0;
}

}  // namespace engine
