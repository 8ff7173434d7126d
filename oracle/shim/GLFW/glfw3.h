// oracle/shim/GLFW/glfw3.h -- TEST INFRASTRUCTURE ONLY.
// src/globals.hpp:22 only needs the opaque window type.
#pragma once
struct GLFWwindow;
struct GLFWmonitor;
