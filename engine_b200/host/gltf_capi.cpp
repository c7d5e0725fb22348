// extern "C" face of gltf_colliders.cpp (include/b2col_gltf.h).
#include <cstring>
#include <new>
#include <string>

#include "../../include/b2col_gltf.h"
#include "gltf_colliders.h"

struct b2col_gltf_scene {
    b2gltf::Scene scene;
    uint32_t n_colliders = 0;
};

namespace {
thread_local std::string g_error;

int Finish(b2gltf::Status st, b2col_gltf_scene* s, b2col_gltf_scene** out)
{
    if (st != b2gltf::OK) {
        delete s;
        return static_cast<int>(st);
    }
    for (const auto& e : s->scene.entities) s->n_colliders += e.has_collider ? 1u : 0u;
    *out = s;
    return B2COL_GLTF_OK;
}
}  // namespace

extern "C" {

int b2col_gltf_load_file(const char* path, uint32_t first_entity, uint32_t root_parent, int strict, b2col_gltf_scene** out)
{
    if (!out || !path) {
        g_error = "null argument";
        return B2COL_GLTF_ERR_FORMAT;
    }
    *out = nullptr;
    auto* s = new (std::nothrow) b2col_gltf_scene();
    if (!s) return B2COL_GLTF_ERR_IO;
    return Finish(b2gltf::LoadFile(path, first_entity, root_parent, strict != 0, s->scene, g_error), s, out);
}

int b2col_gltf_load_memory(const void* data, size_t size, const char* stem, uint32_t first_entity, uint32_t root_parent, int strict, b2col_gltf_scene** out)
{
    if (!out || !data) {
        g_error = "null argument";
        return B2COL_GLTF_ERR_FORMAT;
    }
    *out = nullptr;
    auto* s = new (std::nothrow) b2col_gltf_scene();
    if (!s) return B2COL_GLTF_ERR_IO;
    return Finish(b2gltf::LoadMemory(data, size, stem ? stem : "gltf", first_entity, root_parent, strict != 0, s->scene, g_error), s, out);
}

void b2col_gltf_free(b2col_gltf_scene* s) { delete s; }
const char* b2col_gltf_last_error(void) { return g_error.c_str(); }
uint32_t b2col_gltf_entity_count(const b2col_gltf_scene* s) { return s ? static_cast<uint32_t>(s->scene.entities.size()) : 0u; }
uint32_t b2col_gltf_collider_count(const b2col_gltf_scene* s) { return s ? s->n_colliders : 0u; }

void b2col_gltf_get_entities(const b2col_gltf_scene* s, uint32_t* parent, float* pos3, float* rot_xyzw, float* scl3)
{
    if (!s) return;
    size_t i = 0;
    for (const auto& e : s->scene.entities) {
        if (parent) parent[i] = e.parent;
        if (pos3) std::memcpy(pos3 + 3 * i, e.pos, 12);
        if (rot_xyzw) std::memcpy(rot_xyzw + 4 * i, e.rot, 16);
        if (scl3) std::memcpy(scl3 + 3 * i, e.scl, 12);
        ++i;
    }
}

void b2col_gltf_get_colliders(const b2col_gltf_scene* s, uint32_t* entity, float* aabb6)
{
    if (!s) return;
    size_t k = 0, i = 0;
    for (const auto& e : s->scene.entities) {
        if (e.has_collider) {
            if (entity) entity[k] = s->scene.first_entity + static_cast<uint32_t>(i);
            if (aabb6) std::memcpy(aabb6 + 6 * k, e.aabb, 24);
            ++k;
        }
        ++i;
    }
}

const char* b2col_gltf_entity_name(const b2col_gltf_scene* s, uint32_t i)
{
    return (s && i < s->scene.entities.size()) ? s->scene.entities[i].name.c_str() : "";
}

}  // extern "C"
