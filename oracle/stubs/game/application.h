// TEST INFRASTRUCTURE — headless stand-in for the reference's include/application.h, so that the UNMODIFIED
// reference test/src/camera_controller.cpp (the game's collision-response caller) compiles
// without SDL / Vulkan: same accessor names, scripted input instead of a window.  Only the members that file
// touches exist.  The real Application is never constructed on this path (Scene(nullptr) elsewhere in the
// harness); here Scene receives a pointer to this object and only camera_controller.cpp dereferences it.
#pragma once
#include <filesystem>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include <glm/gtc/constants.hpp>
#include <glm/vec3.hpp>

#include "debug_line.h"

namespace engine {
class Window;
class InputManager;
class SceneManager;
class Renderer;

class Application {
   public:
    Application();
    ~Application();
    std::vector<DebugLine> debug_lines{};
    Window* getWindow() { return window_.get(); }
    InputManager* getInputManager() { return input_.get(); }
    SceneManager* getSceneManager() { return scenes_.get(); }
    // the glTF loader (reference src/gltf_loader.cpp) asks for the renderer and for built-in resources; headless
    // there is neither: stubs/gltf/renderer.h is an empty Renderer, resources are null handles
    Renderer* getRenderer() { return nullptr; }
    template <typename T>
    std::shared_ptr<T> getResource(const std::string&) { return {}; }

   private:
    std::unique_ptr<Window> window_;
    std::unique_ptr<InputManager> input_;
    std::unique_ptr<SceneManager> scenes_;
};
}  // namespace engine
