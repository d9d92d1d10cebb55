// ORACLE — TEST INFRASTRUCTURE ONLY.  Stand-in for libs/gl_utils/utils.h (types only).
#pragma once
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <vector>
#include <string>
#include <GL/glew.h>
#include <glm/glm.hpp>
namespace gl { inline void ExitOnGLError(const char*) {} }
