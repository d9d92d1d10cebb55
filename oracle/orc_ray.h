// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// Per-ray set-up shared by every ray-casting shader of the reference.
#pragma once
#include "orc_common.h"

namespace orc {

struct Ray { vec3 Origin; vec3 Dir; };

// ray_marching_1p.comp:93-99 (the same five lines open every *_marching.comp main()):
//   fpos = storePos + 0.5;  VerPos = (fpos/size)*2 - 1;
//   camera_dir = normalize(vec3(VerPos.x*tan*aspect, VerPos.y*tan, -1) * mat3(u_CameraLookAt));
// `v * M` is the row-vector product: result[j] = dot(v, column j of M).  m3 is column-major.
inline vec3 camera_ray_dir(int px, int py, int w, int h, float tan_fovy, float aspect, const float m3[9]) {
  float fx = (float)px + 0.5f, fy = (float)py + 0.5f;
  float vx = (fx / (float)w) * 2.0f - 1.0f;
  float vy = (fy / (float)h) * 2.0f - 1.0f;
  vec3 v(vx * tan_fovy * aspect, vy * tan_fovy, -1.0f);
  vec3 d(dot(v, vec3(m3[0], m3[1], m3[2])), dot(v, vec3(m3[3], m3[4], m3[5])), dot(v, vec3(m3[6], m3[7], m3[8])));
  return normalize(d);
}

// _common_shaders/ray_bbox_intersection.comp:18-32 (IntersectBox)
inline bool intersect_box(const Ray& r, vec3 boxmin, vec3 boxmax, float& tnear, float& tfar) {
  vec3 invR = vec3(1.0f) / r.Dir;
  vec3 tbbmin = invR * (boxmin - r.Origin);
  vec3 tbbmax = invR * (boxmax - r.Origin);
  vec3 tmin = gmin(tbbmin, tbbmax);
  vec3 tmax = gmax(tbbmin, tbbmax);
  tnear = gmax(gmax(tmin.x, tmin.y), tmin.z);
  tfar = gmin(gmin(tmax.x, tmax.y), tmax.z);
  return tfar > tnear;
}

// ray_bbox_intersection.comp:34-52 (box centred at the origin, extent vol_scaled_dim)
inline bool ray_aabb_intersection(vec3 vert_eye, vec3 vert_dir, vec3 vol_scaled_dim, Ray& r, float& rtnear, float& rtfar) {
  vec3 aabbmin = -vol_scaled_dim * 0.5f;
  vec3 aabbmax = vol_scaled_dim * 0.5f;
  r.Origin = vert_eye;
  r.Dir = normalize(vert_dir);
  float tnear, tfar;
  bool hit = intersect_box(r, aabbmin, aabbmax, tnear, tfar);
  tnear = gmax(tnear, 0.0f);
  rtnear = tnear; rtfar = tfar;
  return hit;
}

// ray_bbox_intersection.comp:54-71 (explicit box)
inline bool ray_aabb_intersection(vec3 vert_eye, vec3 vert_dir, vec3 gridmin, vec3 gridmax, Ray& r, float& rtnear, float& rtfar) {
  r.Origin = vert_eye;
  r.Dir = normalize(vert_dir);
  float tnear, tfar;
  bool hit = intersect_box(r, gridmin, gridmax, tnear, tfar);
  tnear = gmax(tnear, 0.0f);
  rtnear = tnear; rtfar = tfar;
  return hit;
}

}  // namespace orc
