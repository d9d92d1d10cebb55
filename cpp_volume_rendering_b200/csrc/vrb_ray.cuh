// Per-ray set-up shared by every ray-casting kernel.  Written with round-to-nearest intrinsics
// (__fmul_rn/__fadd_rn/__fdiv_rn/__fsqrt_rn are never fused or approximated by nvcc) so that ray
// direction, entry/exit distances and therefore the sample lattice are bit-identical to the float32
// evaluation of the reference shader statements (and to oracle/orc_ray.h).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>

namespace vrb {

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
// GLSL min/max semantics: min(x,y) = y < x ? y : x ; max(x,y) = x < y ? y : x
__device__ __forceinline__ float gmin(float x, float y) { return (y < x) ? y : x; }
__device__ __forceinline__ float gmax(float x, float y) { return (x < y) ? y : x; }

__device__ __forceinline__ float dot_rn(f3 a, f3 b) {
  return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}
__device__ __forceinline__ f3 normalize_rn(f3 a) {
  float l = __fsqrt_rn(dot_rn(a, a));
  return mk3(__fdiv_rn(a.x, l), __fdiv_rn(a.y, l), __fdiv_rn(a.z, l));
}

// ray_marching_1p.comp:93-99 — pixel centre -> NDC -> camera direction (row vector * mat3(LookAt)).
__device__ __forceinline__ f3 camera_ray_dir(int px, int py, int w, int h, float tan_fovy, float aspect, const float* m3) {
  float fx = __fadd_rn((float)px, 0.5f), fy = __fadd_rn((float)py, 0.5f);
  float vx = __fadd_rn(__fmul_rn(__fdiv_rn(fx, (float)w), 2.0f), -1.0f);
  float vy = __fadd_rn(__fmul_rn(__fdiv_rn(fy, (float)h), 2.0f), -1.0f);
  f3 v = mk3(__fmul_rn(__fmul_rn(vx, tan_fovy), aspect), __fmul_rn(vy, tan_fovy), -1.0f);
  f3 d = mk3(dot_rn(v, mk3(m3[0], m3[1], m3[2])), dot_rn(v, mk3(m3[3], m3[4], m3[5])), dot_rn(v, mk3(m3[6], m3[7], m3[8])));
  return normalize_rn(d);
}

// _common_shaders/ray_bbox_intersection.comp:18-52 with explicit box [bmin,bmax]; dir is normalised
// again exactly as RayAABBIntersection does (:40).  Returns hit; tnear is clamped to >= 0 (:46).
__device__ __forceinline__ bool ray_aabb(f3 eye, f3 dir_in, f3 bmin, f3 bmax, f3& dir, float& tnear, float& tfar) {
  dir = normalize_rn(dir_in);
  float ix = __fdiv_rn(1.0f, dir.x), iy = __fdiv_rn(1.0f, dir.y), iz = __fdiv_rn(1.0f, dir.z);
  float ax = __fmul_rn(ix, __fsub_rn(bmin.x, eye.x)), bx = __fmul_rn(ix, __fsub_rn(bmax.x, eye.x));
  float ay = __fmul_rn(iy, __fsub_rn(bmin.y, eye.y)), by = __fmul_rn(iy, __fsub_rn(bmax.y, eye.y));
  float az = __fmul_rn(iz, __fsub_rn(bmin.z, eye.z)), bz = __fmul_rn(iz, __fsub_rn(bmax.z, eye.z));
  float tn = gmax(gmax(gmin(ax, bx), gmin(ay, by)), gmin(az, bz));
  float tf = gmin(gmin(gmax(ax, bx), gmax(ay, by)), gmax(az, bz));
  bool hit = tf > tn;
  tnear = gmax(tn, 0.0f);
  tfar = tf;
  return hit;
}

__device__ __forceinline__ unsigned pack_half2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<unsigned*>(&h);
}

}  // namespace vrb

// ---------------------------------------------------------------------------------------------
// GLSL-statement arithmetic, one IEEE rounding per operation and never contracted: used where a comparison
// downstream (iso crossing, block index, step choice) must come out exactly as in the float32 evaluation of the
// shader statement (and oracle/orc_common.h, built with -ffp-contract=off).
namespace vrb {
__device__ __forceinline__ f3 add3(f3 a, f3 b) { return mk3(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)); }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return mk3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)); }
__device__ __forceinline__ f3 mul3(f3 a, f3 b) { return mk3(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y), __fmul_rn(a.z, b.z)); }
__device__ __forceinline__ f3 div3(f3 a, f3 b) { return mk3(__fdiv_rn(a.x, b.x), __fdiv_rn(a.y, b.y), __fdiv_rn(a.z, b.z)); }
__device__ __forceinline__ f3 scale3(f3 a, float s) { return mk3(__fmul_rn(a.x, s), __fmul_rn(a.y, s), __fmul_rn(a.z, s)); }
__device__ __forceinline__ f3 min3(f3 a, f3 b) { return mk3(gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)); }
__device__ __forceinline__ f3 max3(f3 a, f3 b) { return mk3(gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)); }

// texture(TexVolume, p) on the R16F volume, GL_LINEAR / CLAMP_TO_EDGE with float32 weights, bit-identical to
// oracle Tex3D::linear: two point fetches of the footprint texture (Volume::tex_quad) give the 2x2 texel quads of
// slices i0z and i1z; below texel 0 GL clamps both i0 and i1 to 0, which a zero weight reproduces exactly.
// off: texel origin of a region (bricked) volume inside the whole one (integers; 0 for a whole volume): the texel indices are those of
// the whole volume, the stored array starts at `off`.  A region holds every texel its samples touch (ghost layer), and where it ends at
// the volume's own boundary the array's clamp addressing is the volume's CLAMP_TO_EDGE.
__device__ __forceinline__ float tex3d_f32_exact(cudaTextureObject_t quad, f3 p, f3 res, f3 off = f3{0.0f, 0.0f, 0.0f}) {
  const float X = __fsub_rn(__fmul_rn(p.x, res.x), 0.5f), Y = __fsub_rn(__fmul_rn(p.y, res.y), 0.5f);
  const float Z = __fsub_rn(__fmul_rn(p.z, res.z), 0.5f);
  float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
  const float ax = fx < 0.0f ? 0.0f : __fsub_rn(X, fx), ay = fy < 0.0f ? 0.0f : __fsub_rn(Y, fy), az = __fsub_rn(Z, fz);
  // clamp before the +0.5 so that far-away coordinates stay finite texel addresses
  fx = fminf(fmaxf(fx, -1.0f), res.x) - off.x; fy = fminf(fmaxf(fy, -1.0f), res.y) - off.y; fz = fminf(fmaxf(fz, -1.0f), res.z) - off.z;
  const uint2 q0 = tex3D<uint2>(quad, fx + 0.5f, fy + 0.5f, fz + 0.5f);
  const uint2 q1 = tex3D<uint2>(quad, fx + 0.5f, fy + 0.5f, fz + 1.5f);
  const float2 a0 = __half22float2(*reinterpret_cast<const __half2*>(&q0.x)), b0 = __half22float2(*reinterpret_cast<const __half2*>(&q0.y));
  const float2 a1 = __half22float2(*reinterpret_cast<const __half2*>(&q1.x)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&q1.y));
  const float c00 = __fadd_rn(a0.x, __fmul_rn(ax, __fsub_rn(a0.y, a0.x))), c10 = __fadd_rn(b0.x, __fmul_rn(ax, __fsub_rn(b0.y, b0.x)));
  const float c01 = __fadd_rn(a1.x, __fmul_rn(ax, __fsub_rn(a1.y, a1.x))), c11 = __fadd_rn(b1.x, __fmul_rn(ax, __fsub_rn(b1.y, b1.x)));
  const float c0 = __fadd_rn(c00, __fmul_rn(ay, __fsub_rn(c10, c00))), c1 = __fadd_rn(c01, __fmul_rn(ay, __fsub_rn(c11, c01)));
  return __fadd_rn(c0, __fmul_rn(az, __fsub_rn(c1, c0)));
}

// texture(TexVolumeGradient, p).xyz on the RGBA16F gradient texture, GL_LINEAR / CLAMP_TO_EDGE with float32 weights in the
// order of oracle sample_grad: eight point fetches (texel coordinates, clamp addressing) of Volume::tex_grad_pt.
__device__ __forceinline__ f3 tex3d_rgb_exact(cudaTextureObject_t pt, f3 p, f3 res) {
  const float X = __fsub_rn(__fmul_rn(p.x, res.x), 0.5f), Y = __fsub_rn(__fmul_rn(p.y, res.y), 0.5f), Z = __fsub_rn(__fmul_rn(p.z, res.z), 0.5f);
  float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
  const float ax = __fsub_rn(X, fx), ay = __fsub_rn(Y, fy), az = __fsub_rn(Z, fz);   // below texel 0 both taps clamp to texel 0
  fx = fminf(fmaxf(fx, -1.0f), res.x); fy = fminf(fmaxf(fy, -1.0f), res.y); fz = fminf(fmaxf(fz, -1.0f), res.z);
  float4 t[2][2][2];
#pragma unroll
  for (int k = 0; k < 2; k++)
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
      for (int i = 0; i < 2; i++) t[k][j][i] = tex3D<float4>(pt, fx + 0.5f + (float)i, fy + 0.5f + (float)j, fz + 0.5f + (float)k);
  auto lerp3 = [&](float v000, float v100, float v010, float v110, float v001, float v101, float v011, float v111) {
    const float c00 = __fadd_rn(v000, __fmul_rn(ax, __fsub_rn(v100, v000))), c10 = __fadd_rn(v010, __fmul_rn(ax, __fsub_rn(v110, v010)));
    const float c01 = __fadd_rn(v001, __fmul_rn(ax, __fsub_rn(v101, v001))), c11 = __fadd_rn(v011, __fmul_rn(ax, __fsub_rn(v111, v011)));
    const float c0 = __fadd_rn(c00, __fmul_rn(ay, __fsub_rn(c10, c00))), c1 = __fadd_rn(c01, __fmul_rn(ay, __fsub_rn(c11, c01)));
    return __fadd_rn(c0, __fmul_rn(az, __fsub_rn(c1, c0)));
  };
  return mk3(lerp3(t[0][0][0].x, t[0][0][1].x, t[0][1][0].x, t[0][1][1].x, t[1][0][0].x, t[1][0][1].x, t[1][1][0].x, t[1][1][1].x),
             lerp3(t[0][0][0].y, t[0][0][1].y, t[0][1][0].y, t[0][1][1].y, t[1][0][0].y, t[1][0][1].y, t[1][1][0].y, t[1][1][1].y),
             lerp3(t[0][0][0].z, t[0][0][1].z, t[0][1][0].z, t[0][1][1].z, t[1][0][0].z, t[1][0][1].z, t[1][1][0].z, t[1][1][1].z));
}

// ApplyPhongShading == 1 branch of the ShadeSample functions of the shadowed DVR shaders (rc1pdosct/ray_bbox_marching.comp:630-649,
// rc1pextbsd/ebs_ray_bbox_marching.comp:524-545, rc1pvctsg/vct_ray_bbox_marching.comp:163-183, _common_shaders/obj_ray_marching.comp:
// 236-255).  kind 0: the DOS cone tracer's formula; kind 1: EBS / VCT (Iocc = 1) / object-space march.  Returns false when the
// branch is off (grad_pt == 0) and the caller applies the plain formula; with a zero gradient L keeps the unshaded colour.
struct PhongParams { cudaTextureObject_t grad_pt; float ks, shininess, ispec[3]; };
static __device__ __noinline__ bool phong_shade_sample(int kind, const PhongParams& ph, f3 tx, f3 sizes, f3 res, f3 light_pos, f3 eye,
                                                float ka, float kd, bool apply_shadow, float Iocc, float Isdw, float4& L) {
  if (!ph.grad_pt) return false;
  const float ks = apply_shadow ? ph.ks : 0.0f;
  f3 g = tex3d_rgb_exact(ph.grad_pt, div3(tx, sizes), res);
  if (g.x != 0.0f || g.y != 0.0f || g.z != 0.0f) {
    const f3 Wpos = sub3(tx, scale3(sizes, 0.5f));
    g = normalize_rn(g);
    const f3 light_direction = normalize_rn(sub3(light_pos, Wpos));
    const f3 eye_direction = normalize_rn(sub3(eye, Wpos));
    const f3 halfway = normalize_rn(add3(eye_direction, light_direction));
    const float dd = gmax(0.0f, dot_rn(g, light_direction)), ds = gmax(0.0f, dot_rn(halfway, g));
    const float pw = powf(ds, ph.shininess);
    const float k = __fdiv_rn(1.0f, __fadd_rn(ka, kd));
    const float rgb[3] = {L.x, L.y, L.z};
    float o[3];
    if (kind == 0) {
      const float sc = __fmul_rn(k, __fadd_rn(__fmul_rn(Iocc, ka), __fmul_rn(__fmul_rn(Isdw, kd), dd)));
      const float sp = __fmul_rn(__fmul_rn(Isdw, ks), pw);
#pragma unroll
      for (int i = 0; i < 3; i++) o[i] = __fadd_rn(__fmul_rn(rgb[i], sc), __fmul_rn(ph.ispec[i], sp));
    } else {
#pragma unroll
      for (int i = 0; i < 3; i++) {
        const float amb = __fmul_rn(__fmul_rn(rgb[i], Iocc), ka);
        const float dif = __fmul_rn(Isdw, __fmul_rn(__fmul_rn(rgb[i], kd), dd));
        const float spe = __fmul_rn(Isdw, __fmul_rn(__fmul_rn(ks, ph.ispec[i]), pw));
        o[i] = __fadd_rn(__fmul_rn(k, __fadd_rn(amb, dif)), spe);
      }
    }
    L.x = o[0]; L.y = o[1]; L.z = o[2];
  }
  return true;
}
}  // namespace vrb
