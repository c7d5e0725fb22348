/* b2col_gltf.h — C ABI of the collider ingestion from glTF binaries (libb2col_gltf.so; host-only C++, no CUDA).
 *
 * SURVEY.md 8(f) row 3.  Replaces, for the collision path only, engine::loadGLTF (reference
 * src/gltf_loader.cpp:101-684): parses a .glb container, walks the default scene's node hierarchy exactly as the
 * reference does (:599-679) and returns the entities it would create — in creation order, with parent links and
 * TRS — and the ColliderComponent AABBs it would attach (POSITION accessor min/max, :576-583,647-669).  No
 * meshes, textures or materials are built, so no graphics device is needed.  The arrays plug straight into
 * b2col_set_hierarchy / b2col_set_trs / b2col_set_colliders (include/b2col.h).
 *
 * Entity numbering: the loaded file occupies IDs first_entity .. first_entity + n_entities - 1 (the reference
 * hands out consecutive ascending IDs, src/scene.cpp:37).  Entity 0 of the returned arrays is the file-root entity
 * the reference creates to rotate glTF's Y-up into the engine's Z-up (:599), parent = `root_parent`.
 */
#ifndef B2COL_GLTF_H
#define B2COL_GLTF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2col_gltf_scene b2col_gltf_scene;

enum {
    B2COL_GLTF_OK = 0,
    B2COL_GLTF_ERR_IO = 1,          /* cannot read the file */
    B2COL_GLTF_ERR_FORMAT = 2,      /* not a glTF 2.0 binary container / malformed JSON / bad indices */
    B2COL_GLTF_ERR_UNSUPPORTED = 3  /* a file the reference loader rejects (strict mode), e.g. a primitive
                                       without NORMAL (:381-384) or TEXCOORD_0 (:409-412) */
};

/* strict != 0: fail on the files the reference fails on (same conditions); strict == 0: take every primitive
 * that has a POSITION accessor with min/max.  first_entity: ID the file-root entity will get (>= 1);
 * root_parent: its parent (0 = scene root).  On failure *out is NULL and b2col_gltf_last_error() describes why. */
int b2col_gltf_load_file(const char* path, uint32_t first_entity, uint32_t root_parent, int strict, b2col_gltf_scene** out);
/* same, from memory; `stem` names the file-root entity (the reference uses the file name without extension, :596) */
int b2col_gltf_load_memory(const void* data, size_t size, const char* stem, uint32_t first_entity, uint32_t root_parent, int strict,
                           b2col_gltf_scene** out);
void b2col_gltf_free(b2col_gltf_scene* s);
const char* b2col_gltf_last_error(void);

uint32_t b2col_gltf_entity_count(const b2col_gltf_scene* s);
uint32_t b2col_gltf_collider_count(const b2col_gltf_scene* s);
/* rows = entities in creation order (row i has ID first_entity + i); parent holds entity IDs */
void b2col_gltf_get_entities(const b2col_gltf_scene* s, uint32_t* parent, float* pos3, float* rot_xyzw, float* scl3);
/* colliders in ascending entity ID: entity IDs + local AABBs (min.xyz, max.xyz) */
void b2col_gltf_get_colliders(const b2col_gltf_scene* s, uint32_t* entity, float* aabb6);
/* tag of entity row i (TransformComponent::tag) */
const char* b2col_gltf_entity_name(const b2col_gltf_scene* s, uint32_t i);

#ifdef __cplusplus
}
#endif
#endif /* B2COL_GLTF_H */
