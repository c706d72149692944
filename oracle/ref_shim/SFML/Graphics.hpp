// Headless stand-in for <SFML/Graphics.hpp>; the shim never includes the reference's
// WindowManager.h / Raytracer.h, so nothing beyond the keyboard stub is required.
#pragma once
#include "Window/Keyboard.hpp"
