#include "gltf_loader_colliders.h"

#include <stdexcept>

#include "../host/gltf_colliders.h"
#include "component_collider.h"
#include "component_transform.h"

namespace engine {

Entity loadGLTFColliders(Scene& scene, const std::string& path, bool isStatic)
{
    b2gltf::Scene parsed;
    std::string error;
    const uint32_t first = scene.next_entity_id_;  // IDs are handed out consecutively (reference src/scene.cpp:37)
    if (b2gltf::LoadFile(path, first, 0, /*strict=*/true, parsed, error) != b2gltf::OK) throw std::runtime_error(error);
    Entity root = 0;
    for (const b2gltf::Entity& e : parsed.entities) {
        const Entity id = scene.CreateEntity(e.name, e.parent, glm::vec3{e.pos[0], e.pos[1], e.pos[2]},
                                             glm::quat{e.rot[3], e.rot[0], e.rot[1], e.rot[2]},  // glm::quat(w, x, y, z)
                                             glm::vec3{e.scl[0], e.scl[1], e.scl[2]});
        if (root == 0) root = id;
        scene.GetTransform(id)->is_static = isStatic;
        if (e.has_collider) {
            auto* collider = scene.AddComponent<ColliderComponent>(id);
            collider->aabb.min = {e.aabb[0], e.aabb[1], e.aabb[2]};
            collider->aabb.max = {e.aabb[3], e.aabb[4], e.aabb[5]};
        }
    }
    return root;
}

}  // namespace engine
