// TEST INFRASTRUCTURE — headless stand-in for the reference's include/renderer.h, so that the UNMODIFIED
// reference src/gltf_loader.cpp compiles and runs without a graphics device: meshes,
// textures and materials become empty handles (gltf_resource_stubs.cpp), entities / transforms / colliders are
// created exactly as the reference creates them.
#pragma once
#include "gfx_device.h"
#include "resource_mesh.h"  // the real application.h brings this in through resource_manager.h

namespace engine {
class Renderer {
   public:
    GFXDevice* GetDevice() { return nullptr; }
};
}  // namespace engine
