// TEST INFRASTRUCTURE — stand-in for the reference's include/window.h: no window, no key is ever pressed.
#pragma once
#include "input_keys.h"
#include "input_mouse.h"

namespace engine {
class Window {
   public:
    bool GetKeyPress(inputs::Key) const { return false; }
    bool GetButtonPress(inputs::MouseButton) const { return false; }
    void ToggleFullscreen() {}
    void SetCloseFlag() {}
};
}  // namespace engine
