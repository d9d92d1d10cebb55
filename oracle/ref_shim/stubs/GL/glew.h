// ORACLE — TEST INFRASTRUCTURE ONLY.  Minimal stand-in for <GL/glew.h> so that the reference's GL-free host
// code (which only mentions GL types/enums in texture-creation helpers) compiles without a graphics stack.
// Nothing here talks to a driver.
#pragma once
typedef int GLint; typedef unsigned int GLuint; typedef unsigned int GLenum; typedef void GLvoid;
typedef float GLfloat; typedef int GLsizei; typedef unsigned char GLubyte; typedef unsigned short GLushort;
#define GL_NEAREST 0x2600
#define GL_LINEAR 0x2601
#define GL_CLAMP_TO_EDGE 0x812F
#define GL_CLAMP_TO_BORDER 0x812D
#define GL_RED 0x1903
#define GL_RG 0x8227
#define GL_RGB 0x1907
#define GL_RGBA 0x1908
#define GL_FLOAT 0x1406
#define GL_UNSIGNED_BYTE 0x1401
#define GL_R16F 0x822D
#define GL_R32F 0x822E
#define GL_RG16F 0x822F
#define GL_RGB16F 0x881B
#define GL_RGBA16F 0x881A
#define GL_RGBA32F 0x8814
#define GL_TEXTURE_3D 0x806F
// glTexImage3D of a float RG / R image: recorded per level (ref_api.cpp reads ref_gl_capture::levels)
#include <vector>
namespace ref_gl_capture {
struct Level { int level, internal, w, h, d, channels; std::vector<float> data; };
inline std::vector<Level>& levels() { static std::vector<Level> v; return v; }
}
inline void glTexImage3D(GLenum, GLint level, GLint internalformat, GLsizei w, GLsizei h, GLsizei d, GLint, GLenum format, GLenum type, const void* data) {
  const int ch = format == GL_RED ? 1 : (format == GL_RG ? 2 : (format == GL_RGB ? 3 : 4));
  ref_gl_capture::Level L{level, internalformat, w, h, d, ch, {}};
  if (type == GL_FLOAT && data) L.data.assign((const float*)data, (const float*)data + (size_t)w * h * d * ch);
  ref_gl_capture::levels().push_back(L);
}
