// TEST INFRASTRUCTURE — headless stand-in for the reference's include/resource_material.h (see renderer.h here).
#pragma once
#include <memory>

#include "resource_texture.h"

namespace engine {
class Shader;
class Material {
   public:
    Material(Renderer*, std::shared_ptr<Shader>) {}
    void setAlbedoTexture(std::shared_ptr<Texture>) {}
    void setNormalTexture(std::shared_ptr<Texture>) {}
    void setOcclusionRoughnessMetallicTexture(std::shared_ptr<Texture>) {}
};
}  // namespace engine
