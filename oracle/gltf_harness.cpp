// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// Runs the UNMODIFIED reference glTF loader (reference src/gltf_loader.cpp, compiled where it lies together with
// the vendored tinygltf / json / stb / mikktspace / weldmesh sources) headless inside the reference Scene, and
// dumps what it created on the collision path: entities in creation order with parent + TRS, and the colliders.
// GPU resources (Mesh / Texture / Material) are the no-op handles of stubs/gltf/.  Oracle for
// engine_b200/shim/gltf_colliders.cpp.
#include <cstring>
#include <exception>
#include <string>

#include "component_collider.h"
#include "component_transform.h"
#include "gltf_loader.h"
#include "gltf_loader_colliders.h"  // engine_b200/shim: the collision-path replacement under test
#include "harness_types.h"

using namespace engine;

static std::string g_gltf_error;

extern "C" {

// loadGLTF (src/gltf_loader.cpp:101-684).  Returns the file-root entity, 0 on error (message: gltfh_last_error).
uint32_t gltfh_load(void* h, const char* path, int is_static)
{
    auto* H = static_cast<Harness*>(h);
    try {
        return loadGLTF(H->scene, path, is_static != 0);
    }
    catch (const std::exception& e) {
        g_gltf_error = e.what();
        return 0;
    }
}

// engine_b200/shim's loadGLTFColliders into the same kind of Scene (no GPU involved: pure host code)
uint32_t gltfh_load_b2(void* h, const char* path, int is_static)
{
    auto* H = static_cast<Harness*>(h);
    try {
        return loadGLTFColliders(H->scene, path, is_static != 0);
    }
    catch (const std::exception& e) {
        g_gltf_error = e.what();
        return 0;
    }
}

const char* gltfh_last_error() { return g_gltf_error.c_str(); }

uint32_t gltfh_entity_count(void* h) { return static_cast<Harness*>(h)->scene.next_entity_id_ - 1u; }

// entities first .. first+count-1: parent, position, rotation (x,y,z,w), scale, is_static; tag into a 64-byte slot
void gltfh_dump_entities(void* h, uint32_t first, uint32_t count, uint32_t* parent, float* pos3, float* rot_xyzw, float* scl3, uint8_t* is_static, char* tag64)
{
    auto* H = static_cast<Harness*>(h);
    for (uint32_t i = 0; i < count; ++i) {
        const auto* t = H->scene.GetComponent<TransformComponent>(first + i);
        parent[i] = t->parent;
        std::memcpy(pos3 + 3 * i, &t->position, 12);
        rot_xyzw[4 * i + 0] = t->rotation.x;
        rot_xyzw[4 * i + 1] = t->rotation.y;
        rot_xyzw[4 * i + 2] = t->rotation.z;
        rot_xyzw[4 * i + 3] = t->rotation.w;
        std::memcpy(scl3 + 3 * i, &t->scale, 12);
        is_static[i] = t->is_static ? 1 : 0;
        std::memset(tag64 + 64 * i, 0, 64);
        std::strncpy(tag64 + 64 * i, t->tag.c_str(), 63);
    }
}

// every entity carrying a ColliderComponent, ascending ID; returns the count (call with nulls to size)
uint32_t gltfh_dump_colliders(void* h, uint32_t* entity, float* aabb6)
{
    auto* H = static_cast<Harness*>(h);
    const uint32_t n = H->scene.next_entity_id_ - 1u;
    uint32_t k = 0;
    for (uint32_t e = 1; e <= n; ++e) {
        const auto* c = H->scene.GetComponent<ColliderComponent>(e);
        if (!c) continue;
        if (entity) entity[k] = e;
        if (aabb6) std::memcpy(aabb6 + 6 * k, &c->aabb, 24);
        ++k;
    }
    return k;
}

}  // extern "C"
