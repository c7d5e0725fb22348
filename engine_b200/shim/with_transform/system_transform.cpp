// Bodies of the drop-in engine::TransformSystem (replaces reference src/system_transform.cpp).
//
// Host cost is what matters here: the GPU propagates a million matrices in ~40 us, so a frame is only as fast as the
// walk over the AoS components in front of it.  The reference touches every TransformComponent through
// Scene::GetComponent (two hash look-ups and a dynamic_cast per entity, include/scene.h:56-69,158-170: ~138 ns per
// entity).  This shim
//   * walks the dense component vector through ONE base pointer (include/ecs.h:28-49),
//   * keeps sorted vectors of its entities, rebuilt only after onComponentInsert (std::set iteration is a pointer chase),
//   * after the first frame touches only the entities that can move — TransformComponent::is_static == false
//     (include/component_transform.h:24) — gathers their TRS straight into page-locked staging and uploads those rows
//     (b2col_stage_trs_rows / b2col_commit_trs_rows: 44 bytes per MOVED entity),
//   * writes world_matrix back only where it can have changed: non-static entities and their descendants, gathered on
//     the device into one copy (b2col_fetch_world_rows),
//   * cuts the two per-entity loops that remain (TRS gather, world_matrix scatter: one cache miss per entity on a 144-byte
//     stride) into slices over a few helper threads (b2col_workers.h).
#include "system_transform.h"

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

static_assert(sizeof(glm::quat) == 16 && sizeof(glm::vec3) == 12 && sizeof(glm::mat4) == 64, "glm layouts the packing relies on");

TransformSystem::TransformSystem(Scene* scene) : System(scene, {typeid(TransformComponent).hash_code()}) { shared_ = B2colAcquire(scene); }

TransformSystem::~TransformSystem() { B2colRelease(m_scene); }

void TransformSystem::Check(int status, const char* what) const
{
    if (status != B2COL_OK) throw std::runtime_error(std::string(what) + ": " + b2col_last_error(shared_->ctx));
}

void TransformSystem::onComponentInsert(Entity entity)
{
    (void)entity;
    ++inserted_;
}

// Everything from scratch: entity lists, all rows packed and uploaded, hierarchy (re)levelled on the library side.
void TransformSystem::FullUpload(TransformComponent* base, uint32_t n_entities)
{
    const std::size_t rows = static_cast<std::size_t>(n_entities) + 1u;
    entities_.assign(m_entities.begin(), m_entities.end());  // ascending (std::set order)
    parent_.assign(rows, 0u);
    pos_.assign(3u * rows, 0.0f);
    rot_.assign(4u * rows, 0.0f);
    scl_.assign(3u * rows, 1.0f);
    for (std::size_t e = 0; e < rows; ++e) rot_[4u * e + 3u] = 1.0f;  // rows without a transform stay identity roots
    static_.assign(rows, 0);
    moving_.assign(rows, 0);
    dynamic_.clear();
    writeback_.clear();
    for (Entity e : entities_) {
        const TransformComponent& t = base[e];
        parent_[e] = t.parent;
        std::memcpy(&pos_[3u * e], &t.position, 12);
        std::memcpy(&rot_[4u * e], &t.rotation, 16);  // glm::quat memory order is x,y,z,w
        std::memcpy(&scl_[3u * e], &t.scale, 12);
        static_[e] = t.is_static ? 1 : 0;
        if (!t.is_static) dynamic_.push_back(e);
        // a static child of a moving parent moves in world space (parents have smaller IDs: one ascending pass)
        moving_[e] = (!t.is_static || (t.parent < e && moving_[t.parent])) ? 1 : 0;
        if (moving_[e]) writeback_.push_back(e);
    }
    Check(b2col_set_hierarchy(shared_->ctx, n_entities, parent_.data()), "b2col_set_hierarchy");
    Check(b2col_set_trs(shared_->ctx, n_entities, pos_.data(), rot_.data(), scl_.data()), "b2col_set_trs");
    seen_inserts_ = inserted_;
    uploaded_entities_ = n_entities;
    ++full_uploads;
}

void TransformSystem::onUpdate(float ts)
{
    (void)ts;
    if (m_entities.empty()) return;
    const uint32_t n_entities = m_scene->next_entity_id_ - 1u;
    // The component array is one dense vector indexed by entity ID (reference include/ecs.h:28-49), so one
    // lookup yields the base pointer instead of two hash lookups + a dynamic_cast per entity
    // (include/scene.h:56-69,158-170 — the cost that dominates the reference's TransformSystem).
    const Entity first = *m_entities.begin();
    TransformComponent* const base = m_scene->GetComponent<TransformComponent>(first) - first;

    bool full = !skip_static_rows || seen_inserts_ != inserted_ || uploaded_entities_ != n_entities || entities_.empty();
    if (!full && static_rescan_interval != 0 && ++frames_since_rescan_ >= static_rescan_interval) full = true;  // is_static flags / static rows re-read
    if (!full) {
        // ---- sparse frame: only the entities that may move are read; their rows go straight into page-locked staging
        const uint32_t n = static_cast<uint32_t>(dynamic_.size());
        uint32_t* ids = nullptr;
        float *p = nullptr, *r = nullptr, *s = nullptr;
        Check(b2col_stage_trs_rows(shared_->ctx, n, &ids, &p, &r, &s), "b2col_stage_trs_rows");
        std::atomic<bool> changed{false};  // an entity was re-parented or is no longer dynamic: take the long road this frame
        B2colWorkers::Get().ParallelFor(n, 16384, [&](std::size_t lo, std::size_t hi) {
            for (std::size_t i = lo; i < hi; ++i) {
                const Entity e = dynamic_[i];
                const TransformComponent& t = base[e];
                if (t.parent != parent_[e] || t.is_static) {
                    changed.store(true, std::memory_order_relaxed);
                    return;
                }
                ids[i] = e;
                std::memcpy(p + 3u * i, &t.position, 12);
                std::memcpy(r + 4u * i, &t.rotation, 16);
                std::memcpy(s + 3u * i, &t.scale, 12);
            }
        });
        full = changed.load();
        if (!full) {
            Check(b2col_commit_trs_rows(shared_->ctx, n), "b2col_commit_trs_rows");
            ++sparse_uploads;
        }
        else Check(b2col_commit_trs_rows(shared_->ctx, 0), "b2col_commit_trs_rows");
    }
    const bool first_frame_of_layout = full;
    if (full) {
        FullUpload(base, n_entities);
        frames_since_rescan_ = 0;
    }
    Check(b2col_propagate(shared_->ctx), "b2col_propagate");
    shared_->transforms_on_device = true;
    shared_->n_entities = n_entities;

    // ---- scatter: world_matrix back into the components that the renderer reads
    if (!write_back_world_matrices) return;
    const std::vector<uint32_t>& rows = first_frame_of_layout ? entities_ : writeback_;  // a new layout writes every row once
    const float* w = nullptr;
    Check(b2col_fetch_world_rows(shared_->ctx, static_cast<uint32_t>(rows.size()), rows.data(), &w), "b2col_fetch_world_rows");
    B2colWorkers::Get().ParallelFor(rows.size(), 16384, [&](std::size_t lo, std::size_t hi) {
        for (std::size_t i = lo; i < hi; ++i) std::memcpy(&base[rows[i]].world_matrix, w + 16u * i, 64);
    });
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
