// ORACLE — TEST INFRASTRUCTURE ONLY.  Stand-in for gl::Texture2D: SetData records the float data and formats handed to GL.
#pragma once
#include <gl_utils/utils.h>
namespace gl {
class Texture2D {
 public:
  Texture2D(unsigned int w, unsigned int h) : width(w), height(h) {}
  void GenerateTexture(GLint mn, GLint mg, GLint ws, GLint wt) { min_filter = mn; mag_filter = mg; wrap_s = ws; wrap_t = wt; }
  bool SetData(GLvoid* data, GLint internalformat, GLenum format, GLenum type) {
    internal = internalformat; fmt = format; typ = type;
    const int ch = format == GL_RED ? 1 : (format == GL_RG ? 2 : (format == GL_RGB ? 3 : 4));
    if (type == GL_FLOAT && data) captured.assign((const float*)data, (const float*)data + (size_t)width * height * ch);
    return true;
  }
  unsigned int width, height;
  GLint min_filter = 0, mag_filter = 0, wrap_s = 0, wrap_t = 0, internal = 0;
  GLenum fmt = 0, typ = 0;
  std::vector<float> captured;
};
}  // namespace gl
