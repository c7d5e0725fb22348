// Test-infrastructure stub. reference src/scene.cpp:30 registers a
// MeshRenderSystem whose real translation unit needs Vulkan resources
// (unavailable on this box). The collision path never touches it, so the five
// out-of-line members declared at include/system_mesh_render.h:22-42 are no-ops.
#include "system_mesh_render.h"
#include "component_transform.h"
#include "component_mesh.h"
#include <typeinfo>
namespace engine {
MeshRenderSystem::MeshRenderSystem(Scene* scene)
    : System(scene, {typeid(TransformComponent).hash_code(), typeid(MeshRenderableComponent).hash_code()}) {}
MeshRenderSystem::~MeshRenderSystem() {}
void MeshRenderSystem::RebuildStaticRenderList() {}
void MeshRenderSystem::onComponentInsert(Entity) {}
void MeshRenderSystem::onUpdate(float) {}
}  // namespace engine
