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
#include "../csrc/vrb_synth.cuh"      // the field: one definition shared with the device generator (k_synth.cu), bit for bit

extern "C" {

// dtype: 0 = u8, 1 = u16, 2 = f32.  The region [ox, ox + rw) x [oy, oy + rh) x [oz, oz + rd) of a w x h x d volume, x fastest.
void vrbh_synth_region(int dtype, int w, int h, int d, int ox, int oy, int oz, int rw, int rh, int rd, uint64_t seed, void* out) {
  vrbsynth::Blobs b;
  vrbsynth::make_blobs(seed, b);
#pragma omp parallel for schedule(dynamic, 1)
  for (int z = 0; z < rd; z++)
    for (int y = 0; y < rh; y++) {
      const size_t row = ((size_t)z * rh + y) * (size_t)rw;
      for (int x = 0; x < rw; x++) {
        const float v = vrbsynth::voxel_value(b, dtype, ox + x, oy + y, oz + z, w, h, d);
        if (dtype == 2) ((float*)out)[row + x] = v;
        else if (dtype == 0) ((uint8_t*)out)[row + x] = (uint8_t)v;
        else ((uint16_t*)out)[row + x] = (uint16_t)v;
      }
    }
}

// The z range [z0, z1) of the whole volume (kept for callers that generate slabs).
void vrbh_synth_volume(int dtype, int w, int h, int d, int z0, int z1, uint64_t seed, void* out) {
  vrbh_synth_region(dtype, w, h, d, 0, 0, z0, w, h, z1 - z0, seed, out);
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
