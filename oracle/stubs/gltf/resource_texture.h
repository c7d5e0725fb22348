// TEST INFRASTRUCTURE — headless stand-in for the reference's include/resource_texture.h (see renderer.h here):
// the glTF loader only constructs textures and hands them to materials.
#pragma once
#include <cstdint>

#include "application.h"  // stubs/game
#include "gfx.h"

namespace engine {
class Renderer;
class Texture {
   public:
    Texture(Renderer*, const uint8_t*, int, int, gfx::SamplerInfo, bool) {}
};
}  // namespace engine
