// Collider ingestion from .glb files — C++ face of include/b2col_gltf.h.
// Host-only: the hierarchy walk of the reference's glTF loader without any GPU resource.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace b2gltf {

struct Entity {
    std::string name;   // TransformComponent::tag
    uint32_t parent;    // entity ID (0 = scene root)
    float pos[3];
    float rot[4];       // x, y, z, w
    float scl[3];
    bool has_collider;
    float aabb[6];      // min.xyz, max.xyz — ColliderComponent::aabb
};

struct Scene {
    uint32_t first_entity = 1;
    std::vector<Entity> entities;  // creation order; entity i has ID first_entity + i
};

enum Status { OK = 0, ERR_IO = 1, ERR_FORMAT = 2, ERR_UNSUPPORTED = 3 };

// Returns OK or an error (message in `error`).  See include/b2col_gltf.h for the parameters.
Status LoadMemory(const void* data, size_t size, const std::string& stem, uint32_t first_entity, uint32_t root_parent, bool strict, Scene& out,
                  std::string& error);
Status LoadFile(const std::string& path, uint32_t first_entity, uint32_t root_parent, bool strict, Scene& out, std::string& error);

}  // namespace b2gltf
