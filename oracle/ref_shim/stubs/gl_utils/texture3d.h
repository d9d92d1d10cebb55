// ORACLE — TEST INFRASTRUCTURE ONLY.  Stand-in for gl::Texture3D: no driver; glTexImage3D (GL/glew.h stub) records the data
// each mip level would be uploaded with, so a test can read what the reference's host code hands to GL.
#pragma once
#include <gl_utils/utils.h>
namespace gl {
class Texture3D {
 public:
  Texture3D(unsigned int w, unsigned int h, unsigned int d) : width(w), height(h), depth(d) {}
  void GenerateTexture(GLint, GLint, GLint, GLint, GLint, bool = false) {}
  unsigned int GetWidth() const { return width; }
  unsigned int GetHeight() const { return height; }
  unsigned int GetDepth() const { return depth; }
  unsigned int width, height, depth;
};
}  // namespace gl
