// TEST INFRASTRUCTURE — stand-in for the reference's include/scene_manager.h (only SetActiveScene is named).
#pragma once
namespace engine {
class Scene;
class SceneManager {
   public:
    void SetActiveScene(Scene*) {}
};
}  // namespace engine
