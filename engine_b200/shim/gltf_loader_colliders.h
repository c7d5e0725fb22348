// Collision-path replacement for engine::loadGLTF (reference include/gltf_loader.h:9, src/gltf_loader.cpp:101-684)
// — SURVEY.md 8(f) row 3.  Creates the same entities (same order, names, parents, TRS, is_static) and the same
// ColliderComponents in the Scene as the reference loader, without building meshes / textures / materials, so it
// needs no graphics device: for headless servers, tests and tools that only query collisions.
#pragma once
#include <string>

#include "scene.h"

namespace engine {

// Returns the file-root entity like loadGLTF.  Throws std::runtime_error on the files loadGLTF throws on.
Entity loadGLTFColliders(Scene& scene, const std::string& path, bool isStatic = false);

}  // namespace engine
