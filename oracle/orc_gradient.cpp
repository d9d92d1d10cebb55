// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// Gradient textures the DataManager builds for the Blinn-Phong path, restated from
//   libs/volvis_utils/utils.cpp:287-350 (GenerateSobelFeldmanGradientTexture)
//   libs/volvis_utils/utils.cpp:146-285 (GenerateGradientTexture with its defaults: sample size 1, no filter, normalised)
// Doubles on the CPU, narrowed to float and stored RGB16F (USE_16F_INTERNAL_FORMAT, utils.h:15).
#include "oracle.h"
#include "orc_common.h"
#include <vector>
#include <cstdlib>

using namespace orc;

namespace {
struct Vol {
  const void* v; int dtype, w, h, d;
  double N(int x, int y, int z) const {                          // StructuredGridVolume::GetNormalizedSample: 0 outside
    if (x < 0 || y < 0 || z < 0 || x >= w || y >= h || z >= d) return 0.0;
    const size_t i = (size_t)x + ((size_t)y * w) + ((size_t)z * w * h);
    if (dtype == ORC_U8) return (double)((const uint8_t*)v)[i] / (256.0 - 1.0);
    if (dtype == ORC_U16) return (double)((const uint16_t*)v)[i] / (65536.0 - 1.0);
    return (double)((const float*)v)[i] / (1.0);
  }
};
}  // namespace

// model 3 = COMPUTE_SHADER_SOBEL: main() of libs/volvis_utils/shader/_gradient_shading/sobelfeldman_generator.comp:22-66, dispatched by
// DataManager::GenerateGradientWithComputeShader (datamanager.cpp:623-718).  The shader reads the R16F volume TEXTURE at texel centres
// (GetScalarValue :13-20: 0 outside, else exactly the binary16 texel), accumulates in float32 in the written order
// sg = (sg + a * wp) + b * wn with wp = 4 / pow(2, |v1| + |v2|) (4, 2, 1: exact), stores each channel to an R16F image; the host reads
// the three images back and re-uploads them as one RGB16F texture (no further rounding).
static void sobel_compute_shader(const void* voxels, int dtype, int w, int h, int d, float* out_rgb) {
  std::vector<float> tex((size_t)w * h * d);
  orc_volume_r16f(voxels, dtype, (int64_t)w * h * d, tex.data());
  auto G = [&](int px, int py, int pz) -> float {
    if (px < 0 || py < 0 || pz < 0 || px > w - 1 || py > h - 1 || pz > d - 1) return 0.0f;
    return tex[(size_t)px + ((size_t)py * w) + ((size_t)pz * w * h)];
  };
#pragma omp parallel for collapse(2) schedule(static)
  for (int z = 0; z < d; z++)
    for (int y = 0; y < h; y++)
      for (int x = 0; x < w; x++) {
        float sx = 0.0f, sy = 0.0f, sz = 0.0f;
        for (int v1 = -1; v1 < 2; v1++)
          for (int v2 = -1; v2 < 2; v2++) {
            const int e = std::abs(v1) + std::abs(v2);
            const float wp = e == 0 ? 4.0f : (e == 1 ? 2.0f : 1.0f), wn = -wp;
            sz = (sz + G(x + v1, y + v2, z - 1) * wp) + G(x + v1, y + v2, z + 1) * wn;
            sy = (sy + G(x + v1, y - 1, z + v2) * wp) + G(x + v1, y + 1, z + v2) * wn;
            sx = (sx + G(x - 1, y + v2, z + v1) * wp) + G(x + 1, y + v2, z + v1) * wn;
          }
        float* o = out_rgb + 3 * ((size_t)x + ((size_t)y * w) + ((size_t)z * w * h));
        o[0] = round_f16(sx); o[1] = round_f16(sy); o[2] = round_f16(sz);
      }
}

// model 1 = SOBEL_FELDMAN_FILTER, 2 = FINITE_DIFERENCES, 3 = COMPUTE_SHADER_SOBEL.  out_rgb: w*h*d*3 floats (RGB16F values).
extern "C" void orc_gradient(const void* voxels, int dtype, int w, int h, int d, int model, float* out_rgb) {
  if (model == 3) { sobel_compute_shader(voxels, dtype, w, h, d, out_rgb); return; }
  Vol V{voxels, dtype, w, h, d};
#pragma omp parallel for collapse(2) schedule(static)
  for (int z = 0; z < d; z++)
    for (int y = 0; y < h; y++)
      for (int x = 0; x < w; x++) {
        double gx = 0.0, gy = 0.0, gz = 0.0;
        if (model == 1) {
          for (int v1 = -1; v1 <= 1; v1++)
            for (int v2 = -1; v2 <= 1; v2++) {
              const double wp = 4.0 / std::pow(2.0, std::abs(v1) + std::abs(v2)), wn = -4.0 / std::pow(2.0, std::abs(v1) + std::abs(v2));
              gz += V.N(x + v1, y + v2, z - 1) * wp + V.N(x + v1, y + v2, z + 1) * wn;
              gy += V.N(x + v1, y - 1, z + v2) * wp + V.N(x + v1, y + 1, z + v2) * wn;
              gx += V.N(x - 1, y + v2, z + v1) * wp + V.N(x + 1, y + v2, z + v1) * wn;
            }
        } else {
          gx = V.N(x + 1, y, z) - V.N(x - 1, y, z);
          gy = V.N(x, y + 1, z) - V.N(x, y - 1, z);
          gz = V.N(x, y, z + 1) - V.N(x, y, z - 1);
          const double inv = 1.0 / std::sqrt((gx * gx + gy * gy) + gz * gz);     // glm::normalize = v * inversesqrt(dot(v, v))
          gx = gx * inv; gy = gy * inv; gz = gz * inv;
          if (gx != gx) { gx = 0.0; gy = 0.0; gz = 0.0; }                          // NaN test on x only, as the reference
        }
        float* o = out_rgb + 3 * ((size_t)x + ((size_t)y * w) + ((size_t)z * w * h));
        o[0] = round_f16((float)gx); o[1] = round_f16((float)gy); o[2] = round_f16((float)gz);
      }
}
