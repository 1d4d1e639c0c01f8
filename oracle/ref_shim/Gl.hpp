/*
 * oracle/ref_shim/Gl.hpp -- TEST INFRASTRUCTURE ONLY.
 *
 * Shadows the reference's include/Gl.hpp (glbinding + GLFW) when the reference
 * headers are compiled headless for the oracle.  include/Globals.hpp:5,39-85
 * only needs the GL *type names* to declare viewer globals that the tracing
 * path never touches.  There is no graphics stack on either box.
 */
#pragma once

typedef unsigned int GLuint;
typedef int GLint;
typedef char GLchar;
typedef float GLfloat;
struct GLFWwindow;
