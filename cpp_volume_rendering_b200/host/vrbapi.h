// Run-time binding of libvrb200.so (include/vrb200.h) for the host library: dlopen once, resolve every entry point
// the plugins use.  No fallback: if the library or a device is missing, VrbApi::Get() reports why and every
// renderer Init() returns false (the reference's behaviour when a resource is missing, rc1prenderer.cpp:54).
#pragma once
#include <string>
#include "vrb200.h"

struct VrbApi {
  const char* (*last_error)(void);
  int (*ctx_create)(int, void*, vrb_ctx**);
  int (*ctx_destroy)(vrb_ctx*);
  int (*ctx_synchronize)(vrb_ctx*);
  int (*volume_upload)(vrb_ctx*, const void*, int, int64_t, int64_t, int64_t, float, float, float);
  int (*tf_upload)(vrb_ctx*, const double*, int, int);
  int (*volume_gradient)(vrb_ctx*, int, float*);
  int (*renderer_create)(vrb_ctx*, int, vrb_renderer**);
  int (*renderer_destroy)(vrb_renderer*);
  int (*renderer_init)(vrb_renderer*, int, int);
  int (*renderer_reshape)(vrb_renderer*, int, int);
  int (*renderer_set_params)(vrb_renderer*, const void*, size_t);
  int (*renderer_get_params)(vrb_renderer*, void*, size_t);
  int (*renderer_update)(vrb_renderer*, const vrb_camera*, const vrb_light*);
  int (*renderer_redraw)(vrb_renderer*, int);
  int (*renderer_set_multiscaling)(vrb_renderer*, int, int);
  int (*crtgt_set_kernels)(vrb_renderer*, const float*, int, const float*, int);
  int (*crtgt_read_state)(vrb_renderer*, float*, long long*);
  int (*renderer_frame)(vrb_renderer*, void**, int*, int*);
  int (*renderer_read_frame)(vrb_renderer*, void*, int);
  int (*renderer_counters)(vrb_renderer*, uint64_t*, double*);
  // nullptr + message when libvrb200.so cannot be loaded
  static const VrbApi* Get(std::string* why = nullptr);
};
