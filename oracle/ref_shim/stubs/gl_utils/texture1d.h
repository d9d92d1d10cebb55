// ORACLE — TEST INFRASTRUCTURE ONLY.  Capturing stand-in for gl::Texture1D (libs/gl_utils/texture1d.h:23-43):
// same public interface; SetData keeps the float data the reference hands to glTexImage1D so that the
// reference's own GenerateTexture_1D_RGBt/_RGBA output can be read back by the tests.
#pragma once
#include <gl_utils/utils.h>
namespace gl {
class Texture1D {
 public:
  explicit Texture1D(unsigned int length) : m_length(length) {}
  ~Texture1D() {}
  void GenerateTexture(GLint mn, GLint mx, GLint wrap) { min_filter = mn; mag_filter = mx; wrap_s = wrap; }
  bool SetData(GLvoid* data, GLint internalformat, GLenum format, GLenum type) {
    int ch = format == GL_RGBA ? 4 : (format == GL_RGB ? 3 : (format == GL_RG ? 2 : 1));
    if (type != GL_FLOAT) return false;
    captured.assign((float*)data, (float*)data + (size_t)m_length * ch);
    internal = internalformat; channels = ch;
    return true;
  }
  GLuint GetTextureID() { return 0; }
  unsigned int GetLength() { return m_length; }
  std::vector<float> captured; GLint internal = 0, min_filter = 0, mag_filter = 0, wrap_s = 0; int channels = 0;
 private:
  unsigned int m_length;
};
}
