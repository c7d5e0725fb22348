// TEST INFRASTRUCTURE — state shared by ref_harness.cpp and game_harness.cpp (one handle, two C APIs).
#pragma once
#include <vector>

#include "scene.h"

namespace engine {
class Application;
}
// stubs/game/application_stub.cpp: the headless Application stand-in (only the game's camera controller
// dereferences Scene::app(); nothing on the collision path does)
engine::Application* b2h_make_app();
void b2h_free_app(engine::Application*);

struct Harness {
    engine::Application* app;
    engine::Scene scene;
    std::vector<engine::Entity> colliders;
    engine::Entity camera = 0;
    Harness() : app(b2h_make_app()), scene(app) {}
    ~Harness() { b2h_free_app(app); }
    Harness(const Harness&) = delete;
    Harness& operator=(const Harness&) = delete;
};
