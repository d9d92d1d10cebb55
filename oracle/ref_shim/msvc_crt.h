// ORACLE — test infrastructure only.  The MSVC CRT names libs/file_utils/pvm.cpp uses, for g++ (force-included by the
// Makefile when the reference's sources are compiled where they lie).
#pragma once
#include <cstdio>
#ifndef _MSC_VER
#define sscanf_s sscanf
typedef int errno_t;
static inline int fopen_s(FILE** f, const char* name, const char* mode) { *f = std::fopen(name, mode); return *f ? 0 : 1; }
#ifdef __cplusplus
#include <cmath>
using std::sin; using std::cos;   // MSVC's <cmath> puts the float overloads in the global namespace (camera.cpp: sin(float) * glm::vec3)
#endif
#endif
