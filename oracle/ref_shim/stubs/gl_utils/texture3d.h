// ORACLE — TEST INFRASTRUCTURE ONLY.  Stand-in for gl::Texture3D (declaration only; unused by the SAT code).
#pragma once
#include <gl_utils/utils.h>
namespace gl { class Texture3D; }
