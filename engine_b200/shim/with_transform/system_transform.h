// Drop-in replacement for the reference's include/system_transform.h (class engine::TransformSystem,
// /root/reference/include/system_transform.h:7-21): same namespace, class name, base class and methods.
// onUpdate packs the AoS TransformComponents (144 B each, include/component_transform.h:13-25) into SoA
// staging buffers, runs the propagation on the GPU (b2col_propagate) and writes world_matrix back for the
// consumers that stay on the CPU (MeshRenderSystem, src/system_mesh_render.cpp:62).  The matrices stay on the
// device for the drop-in CollisionSystem of the same Scene.
#pragma once

#include <string>
#include <vector>

#include "ecs.h"

struct b2col_ctx;

namespace engine {

struct B2colSceneCtx;

class TransformSystem : public System {
   public:
    explicit TransformSystem(Scene* scene);
    ~TransformSystem() override;

    void onUpdate(float ts) override;

    // Linear search among `parent`'s children for `tag` (reference src/system_transform.cpp:40-51).
    Entity GetChildEntity(Entity parent, const std::string& tag);

    // Static subtrees (TransformComponent::is_static, include/component_transform.h:24) are uploaded once:
    // while nothing but non-static TRS values change, only those rows are re-packed.
    bool skip_static_rows = true;

   private:
    B2colSceneCtx* shared_ = nullptr;
    std::vector<uint32_t> parent_, parent_prev_;
    std::vector<float> pos_, rot_, scl_, world_;
    std::vector<unsigned char> static_;
    bool packed_once_ = false;
    void Check(int status, const char* what) const;
};

}  // namespace engine
