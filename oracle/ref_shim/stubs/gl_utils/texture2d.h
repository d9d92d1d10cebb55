// ORACLE — TEST INFRASTRUCTURE ONLY.  Stand-in for gl::Texture2D (declaration only; unused by the SAT code).
#pragma once
#include <gl_utils/utils.h>
namespace gl { class Texture2D; }
