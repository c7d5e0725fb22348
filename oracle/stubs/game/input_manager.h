// TEST INFRASTRUCTURE — scripted stand-in for the reference's include/input_manager.h (see application.h here).
#pragma once
#include <string>

namespace engine {
class InputManager {
   public:
    // the frame's scripted input, set by the harness before Scene::Update
    float movex = 0.0f, movey = 0.0f, lookx = 0.0f, looky = 0.0f;
    bool sprint = false, jump = false;
    float GetAxis(const std::string& axis_name) const
    {
        if (axis_name == "movex") return movex;
        if (axis_name == "movey") return movey;
        if (axis_name == "lookx") return lookx;
        if (axis_name == "looky") return looky;
        return 0.0f;
    }
    bool GetButton(const std::string& button_name) const { return button_name == "sprint" ? sprint : false; }
    bool GetButtonPress(const std::string& button_name) const { return button_name == "jump" ? jump : false; }
};
}  // namespace engine
