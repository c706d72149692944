// Headless stand-in for <SFML/Window/Keyboard.hpp> (SFML 2.6 is not vendored upstream and
// there is no window system on the build or GPU boxes). Only what
// CudaRayTracer/src/Core/Camera.cuh:52,140-151 names. No key is ever pressed.
#pragma once
namespace sf {
class Keyboard {
public:
    enum Key { A, D, F, S, W, Space, LControl, PageUp, PageDown };
    static bool isKeyPressed(Key) { return false; }
};
}  // namespace sf
