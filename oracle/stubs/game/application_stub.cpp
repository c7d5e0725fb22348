// TEST INFRASTRUCTURE — out-of-line members of the headless Application stand-in.
#include "application.h"
#include "input_manager.h"
#include "scene_manager.h"
#include "window.h"

namespace engine {
Application::Application() : window_(new Window()), input_(new InputManager()), scenes_(new SceneManager()) {}
Application::~Application() = default;
}  // namespace engine

engine::Application* b2h_make_app() { return new engine::Application(); }
void b2h_free_app(engine::Application* a) { delete a; }
