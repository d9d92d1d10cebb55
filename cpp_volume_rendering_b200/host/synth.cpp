// Deterministic synthetic volumes for parity tests and benchmarks (no reference
// counterpart — the reference ships no dataset in this checkout).
//   f(p) = max_k A_k exp(-|p-c_k|^2 / 2 s_k^2) + 0.15 [ | |p-0.5| - 0.33 | < 0.01 ],  p = voxel centre in [0,1]^3
// K = 24 blobs drawn from SplitMix64(seed); +-1 LSB value noise from a counter hash; quantised to
// u8 = round(255 f), u16 = round(65535 f) or stored as f32.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <vector>
#include <algorithm>

namespace {
struct SplitMix64 {
  uint64_t s;
  explicit SplitMix64(uint64_t seed) : s(seed) {}
  uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
  double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};
inline uint32_t hash3(uint32_t x, uint32_t y, uint32_t z, uint32_t seed) {
  uint32_t h = seed ^ (x * 0x9E3779B1u) ^ (y * 0x85EBCA77u) ^ (z * 0xC2B2AE3Du);
  h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
  return h;
}
constexpr int K = 24;
}  // namespace

extern "C" {

// dtype: 0 = u8, 1 = u16, 2 = f32.  out holds w*h*d elements, x fastest.  z range [z0,z1) lets the
// caller generate a brick without materialising the whole volume (2048^3 case).
void vrbh_synth_volume(int dtype, int w, int h, int d, int z0, int z1, uint64_t seed, void* out) {
  float cx[K], cy[K], cz[K], inv2s2[K], amp[K], reach2[K];
  SplitMix64 rng(seed);
  for (int k = 0; k < K; k++) {
    cx[k] = (float)(0.1 + 0.8 * rng.uniform()); cy[k] = (float)(0.1 + 0.8 * rng.uniform()); cz[k] = (float)(0.1 + 0.8 * rng.uniform());
    float s = (float)(0.03 + 0.09 * rng.uniform());
    amp[k] = (float)(0.3 + 0.7 * rng.uniform());
    inv2s2[k] = 1.0f / (2.0f * s * s);
    reach2[k] = 16.0f * s * s * 2.0f;  // beyond ~5.6 sigma the blob is < 1e-7: skip the exp
  }
  const uint32_t nseed = (uint32_t)(seed ^ 0x5bd1e995u) + 1u;
  const float maxq = dtype == 0 ? 255.0f : 65535.0f;
#pragma omp parallel for schedule(dynamic, 1)
  for (int z = z0; z < z1; z++) {
    const float pz = ((float)z + 0.5f) / (float)d;
    for (int y = 0; y < h; y++) {
      const float py = ((float)y + 0.5f) / (float)h;
      size_t row = ((size_t)(z - z0) * h + y) * (size_t)w;
      for (int x = 0; x < w; x++) {
        const float px = ((float)x + 0.5f) / (float)w;
        float f = 0.0f;
        for (int k = 0; k < K; k++) {
          float dx = px - cx[k], dy = py - cy[k], dz = pz - cz[k];
          float r2 = dx * dx + dy * dy + dz * dz;
          if (r2 < reach2[k]) f = std::max(f, amp[k] * std::exp(-r2 * inv2s2[k]));
        }
        float ex = px - 0.5f, ey = py - 0.5f, ez = pz - 0.5f;
        float rr = std::sqrt(ex * ex + ey * ey + ez * ez);
        if (std::fabs(rr - 0.33f) < 0.01f) f += 0.15f;
        f = std::min(std::max(f, 0.0f), 1.0f);
        int noise = (int)(hash3((uint32_t)x, (uint32_t)y, (uint32_t)z, nseed) % 3u) - 1;
        if (dtype == 2) {
          ((float*)out)[row + x] = std::min(std::max(f + (float)noise * (1.0f / 65535.0f), 0.0f), 1.0f);
        } else {
          float q = std::floor(f * maxq + 0.5f) + (float)noise;
          q = std::min(std::max(q, 0.0f), maxq);
          if (dtype == 0) ((uint8_t*)out)[row + x] = (uint8_t)q; else ((uint16_t*)out)[row + x] = (uint16_t)q;
        }
      }
    }
  }
}

// Writes an 8-bit volume as a .syn text file in the grammar of VolumeReader::readsyn (libs/volvis_utils/reader.cpp:283-371): "W H D",
// a first box record "1 0 0 0 W H D 0" that defines the buffer the reference leaves uninitialised, then one record per run of equal
// non-zero voxels along x — "1 x0 y z x1 y+1 z+1 v" for a run, "0 x y z v" for a single voxel.  Returns the number of records, -1 on
// an I/O error.  (The 256^3 volume of BASELINE config 1 has ~10^7 runs: this is the same writer as inputs.write_syn, without Python's
// per-record cost.)
long long vrbh_write_syn(const char* path, const uint8_t* vox, int w, int h, int d) {
  FILE* f = std::fopen(path, "w");
  if (!f) return -1;
  std::vector<char> buf(1 << 22);
  std::setvbuf(f, buf.data(), _IOFBF, buf.size());
  long long records = 1;
  std::fprintf(f, "%d %d %d\n1 0 0 0 %d %d %d 0\n", w, h, d, w, h, d);
  for (int z = 0; z < d; z++)
    for (int y = 0; y < h; y++) {
      const uint8_t* row = vox + ((size_t)z * h + y) * (size_t)w;
      int x0 = 0;
      while (x0 < w) {
        int x1 = x0 + 1;
        while (x1 < w && row[x1] == row[x0]) x1++;
        if (row[x0]) {
          if (x1 - x0 > 1) std::fprintf(f, "1 %d %d %d %d %d %d %d\n", x0, y, z, x1, y + 1, z + 1, (int)row[x0]);
          else std::fprintf(f, "0 %d %d %d %d\n", x0, y, z, (int)row[x0]);
          records++;
        }
        x0 = x1;
      }
    }
  const bool ok = std::fflush(f) == 0 && !std::ferror(f);
  std::fclose(f);
  return ok ? records : -1;
}

}  // extern "C"
