// Drop-in replacement for the reference's include/system_transform.h (class engine::TransformSystem,
// /root/reference/include/system_transform.h:7-21): same namespace, class name, base class and methods.
// onUpdate packs the AoS TransformComponents (144 B each, include/component_transform.h:13-25) into SoA
// staging buffers, runs the propagation on the GPU (b2col_propagate) and writes world_matrix back for the
// consumers that stay on the CPU (MeshRenderSystem, src/system_mesh_render.cpp:62).  The matrices stay on the
// device for the drop-in CollisionSystem of the same Scene.
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "ecs.h"

struct b2col_ctx;

namespace engine {

struct B2colSceneCtx;
struct TransformComponent;

class TransformSystem : public System {
   public:
    explicit TransformSystem(Scene* scene);
    ~TransformSystem() override;

    void onUpdate(float ts) override;
    void onComponentInsert(Entity entity) override;

    // Linear search among `parent`'s children for `tag` (reference src/system_transform.cpp:40-51).
    Entity GetChildEntity(Entity parent, const std::string& tag);

    // Static subtrees (TransformComponent::is_static, include/component_transform.h:24) are packed and uploaded once: after the
    // first frame of an entity layout only the non-static entities are read (their TRS rows are uploaded sparsely) and only they
    // and their descendants get world_matrix written back.  is_static is taken as a promise: a static entity whose TRS the game
    // changes anyway is picked up at the next full upload — a new entity, a re-parented or newly static dynamic entity, or
    // every static_rescan_interval frames when that is non-zero.  skip_static_rows = false restores the reference's behaviour
    // of reading every entity every frame.
    bool skip_static_rows = true;
    unsigned static_rescan_interval = 0;
    // world_matrix is only read on the CPU by MeshRenderSystem (src/system_mesh_render.cpp:62); a headless server can switch the
    // write-back off (the matrices stay on the device for the collision system either way)
    bool write_back_world_matrices = true;
    // statistics (tests, bench)
    unsigned long full_uploads = 0, sparse_uploads = 0;

   private:
    B2colSceneCtx* shared_ = nullptr;
    std::vector<uint32_t> entities_, dynamic_, writeback_;  // ascending IDs: all / is_static == false / those whose world matrix can change
    std::vector<uint32_t> parent_;
    std::vector<float> pos_, rot_, scl_;
    std::vector<unsigned char> static_, moving_;
    std::size_t inserted_ = 0, seen_inserts_ = ~std::size_t(0);
    uint32_t uploaded_entities_ = 0;
    unsigned frames_since_rescan_ = 0;
    void FullUpload(TransformComponent* base, uint32_t n_entities);
    void Check(int status, const char* what) const;
};

}  // namespace engine
