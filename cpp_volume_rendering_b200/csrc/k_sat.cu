// K12 — the extinction summed-area table of RC1PExtinctionBasedShading, moved from the CPU to the GPU.
//   reference: GenerateExtinctionSAT3DTex, cppvolrend/structured/rc1pextbsd/ebsrenderer.cpp:624-722 (fill, cast, upload)
//              SummedAreaTable3D<double>::BuildSAT, libs/vis_utils/summedareatable.h:218-278 (serial CPU recurrence)
//
// (W+2)(H+2)(D+2) cells, zero border, interior = TransferFunction1D::GetExtN(normalised voxel) (a float, widened to
// double), inclusive 3D prefix sum accumulated in double, stored as float32.
//
// Three builders (vrb_ebs_params::fast_sat = 0 / 1 / 2):
//  * EXACT (default): the reference's own 8-term inclusion-exclusion recurrence, term for term, evaluated as a
//    wavefront over the planes x+y+z = k (every cell depends only on cells of earlier planes).  The double results —
//    and so the float32 table — are bit-identical to the CPU build for any input.
//  * FAST (1): the reduce-then-scan of `namespace sat2` below when the volume's x extent allows it, else the separable scan.
//  * SEPARABLE (2): separable scan (x by warp shuffles, then y, then z fused with the float32 store).  HBM-bound: 56 B/cell
//    moved against 8 B/cell algorithmic.  Re-associates the double sums: identical after the float32 cast except for
//    rare 1-ulp cells, and bit-identical whenever the sums are exact in double (quantised extinctions).
#include "vrb_internal.cuh"
#include "vrb_sat.cuh"
#include <cmath>
#include <vector>

namespace vrb {

// TransferFunction1D::Get(v, 1.0).a -> GetExtN (transferfunction1d.cpp:132-157,189-197; transferfunction.h:79-82)
__host__ __device__ inline float tf_get_extn(const double* table, int max_density, int extinction_type, double nv) {
  double value = nv * ((double)max_density / 1.0);
  float a;
  if (value < 0.0f || value > (float)max_density) a = 0.0f;
  else if (fabs(value - (float)max_density) < 0.000001) a = (float)table[4 * max_density + 3];
  else {
    const int iv = (int)value;
    const double t = value - iv;
    a = (float)((1.0 - t) * table[4 * iv + 3] + t * table[4 * (iv + 1) + 3]);
  }
  if (!extinction_type) return a == 0.0f ? 0.0f : (float)log(1.0 / (1.0 - (double)a));   // log(1/1) = 0 exactly: skip the double log in empty space
  return a;
}

// Cell (x,y,z) of the bordered grid <- extinction.  Integer volumes read a per-value table built on the host with
// the host's libm (the same arithmetic the reference runs); float volumes evaluate GetExtN on the device.
template <typename T>
__global__ void __launch_bounds__(256) sat_fill_kernel(const T* __restrict__ raw, const float* __restrict__ lut,
                                                       const double* __restrict__ table, int max_density, int ext_type,
                                                       int w, int h, int d, double* __restrict__ acc) {
  const int W = w + 2, H = h + 2;
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= W) return;
  double v = 0.0;
  if (x >= 1 && x <= w && y >= 1 && y <= h && z >= 1 && z <= d) {
    const size_t i = (size_t)(x - 1) + (size_t)w * ((size_t)(y - 1) + (size_t)h * (size_t)(z - 1));
    if (lut) v = (double)lut[(unsigned)raw[i]];
    else v = (double)tf_get_extn(table, max_density, ext_type, (double)raw[i] / 1.0);
  }
  acc[(size_t)x + (size_t)W * ((size_t)y + (size_t)H * (size_t)z)] = v;
}

// ---- EXACT: BuildSAT (summedareatable.h:218-278) as a wavefront over x + y + z, in place ----------------------------------------------
// Step 1 (the origin), step 2 (the three axes: two terms), step 3 (the three faces: four terms), step 4 (the interior: eight terms), each
// with the operand order of the reference's statement — the double sums round at every cell, so the order is part of the result.
// (r1 launched one kernel per plane of cells: 1539 launches, 16 ms at 514^3.)
// Tile by tile: a CTA owns a 32 x 32 x 8 block of cells, loads it with the faces / edges / corner of the three blocks
// behind it into shared memory, runs the cell wavefront x + y + z = s there (70 barriers), and writes the block back; the blocks themselves
// are launched diagonal by diagonal (97 launches at 514^3 instead of 1539, each with coalesced global traffic).  Every cell is computed by
// the statement its position selects, from the same operands as the serial loop: the same bits.
constexpr int WT_X = 32, WT_Y = 32, WT_Z = 8, WT_SX = 34, WT_SY = WT_Y + 1;       // row stride 34 doubles: a diagonal's cells fall into different banks
__global__ void __launch_bounds__(256) sat_tile_wavefront_kernel(double* __restrict__ a, int W, int H, int D, int kd, int ntx, int ntz) {
  const int tx = blockIdx.x % ntx, ty = blockIdx.x / ntx, tz = kd - tx - ty;
  if (tz < 0 || tz >= ntz) return;
  extern __shared__ double wt_sm[];                                                 // [WT_Z + 1][WT_Y + 1][WT_SX], index 0 = the cell behind the block
  const int x0 = tx * WT_X, y0 = ty * WT_Y, z0 = tz * WT_Z;
  const size_t sy = (size_t)W, sz = (size_t)W * H;
  for (int i = threadIdx.x; i < (WT_Z + 1) * WT_SY * (WT_X + 1); i += blockDim.x) {
    const int ix = i % (WT_X + 1), iy = (i / (WT_X + 1)) % WT_SY, iz = i / ((WT_X + 1) * WT_SY);
    const int gx = x0 - 1 + ix, gy = y0 - 1 + iy, gz = z0 - 1 + iz;
    double v = 0.0;
    if (gx >= 0 && gx < W && gy >= 0 && gy < H && gz >= 0 && gz < D) v = a[(size_t)gx + sy * gy + sz * gz];
    wt_sm[(iz * WT_SY + iy) * WT_SX + ix] = v;
  }
  __syncthreads();
  const int lz = threadIdx.x >> 5, lx = threadIdx.x & 31;
  const int x = x0 + lx, z = z0 + lz;
  for (int s = 0; s < WT_X + WT_Y + WT_Z - 2; s++) {
    const int ly = s - lz - lx, y = y0 + ly;
    if (ly >= 0 && ly < WT_Y && x < W && y < H && z < D) {
      double* c = wt_sm + ((lz + 1) * WT_SY + (ly + 1)) * WT_SX + (lx + 1);
      constexpr int SY = WT_SX, SZ = WT_SY * WT_SX;
      const int nz = (x > 0) + (y > 0) + (z > 0);
      if (nz == 1) c[0] = (x > 0 ? c[-1] : (y > 0 ? c[-SY] : c[-SZ])) + c[0];
      else if (nz == 2) {
        if (y == 0) c[0] = c[-1] + c[-SZ] - c[-1 - SZ] + c[0];
        else if (z == 0) c[0] = c[-1] + c[-SY] - c[-1 - SY] + c[0];
        else c[0] = c[-SY] + c[-SZ] - c[-SY - SZ] + c[0];
      } else if (nz == 3) {
        c[0] = c[0] + c[-1 - SY - SZ] + c[-SZ] + c[-SY] + c[-1] - c[-1 - SY] - c[-SY - SZ] - c[-1 - SZ];
      }
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < WT_Z * WT_Y * WT_X; i += blockDim.x) {
    const int ix = i % WT_X, iy = (i / WT_X) % WT_Y, iz = i / (WT_X * WT_Y);
    const int gx = x0 + ix, gy = y0 + iy, gz = z0 + iz;
    if (gx < W && gy < H && gz < D) a[(size_t)gx + sy * gy + sz * gz] = wt_sm[((iz + 1) * WT_SY + (iy + 1)) * WT_SX + (ix + 1)];
  }
}

__global__ void __launch_bounds__(256) sat_to_float_kernel(const double* __restrict__ a, float* __restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (float)a[i];
}

// ---- FAST: separable scan ---------------------------------------------------------------------------------------
// x: one warp per row, 32 cells per step, Kogge-Stone in double + running carry.
__global__ void __launch_bounds__(256) sat_scan_x_kernel(double* __restrict__ a, int W, size_t rows) {
  const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  double* r = a + row * (size_t)W;
  double carry = 0.0;
  for (int x0 = 0; x0 < W; x0 += 32) {
    const int x = x0 + lane;
    double v = x < W ? r[x] : 0.0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    v += carry;
    if (x < W) r[x] = v;
    carry = __shfl_sync(0xffffffffu, v, 31);
  }
}
// y: one thread per (x, z) column walks y; consecutive threads = consecutive x (coalesced).
__global__ void __launch_bounds__(256) sat_scan_y_kernel(double* __restrict__ a, int W, int H, int D) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, z = blockIdx.y;
  if (x >= W) return;
  double* p = a + (size_t)x + (size_t)W * H * (size_t)z;
  double run = 0.0;
  for (int y = 0; y < H; y++) { run += p[(size_t)y * W]; p[(size_t)y * W] = run; }
}
// z: one thread per (x, y) walks z and stores the float32 table.
__global__ void __launch_bounds__(256) sat_scan_z_kernel(const double* __restrict__ a, float* __restrict__ out, int W, int H, int D) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const size_t base = (size_t)x + (size_t)W * y, sz = (size_t)W * H;
  double run = 0.0;
  for (int z = 0; z < D; z++) { run += a[base + sz * z]; out[base + sz * z] = (float)run; }
}

// ---- FAST v2: reduce-then-scan, the volume is read twice and the table written once -------------------------------
// A CTA of 16 warps owns a tile of 512 (x) x 32 (y) cells — thread = 4 consecutive x, 8 consecutive y — and marches a
// segment of z slices.  Slices are staged global -> shared by cp.async (two stages in flight, every thread reads back
// only the bytes it copied itself, so the staging needs no barrier).
//   pass 1 `sat_reduce_kernel`: extinction E(x,y,z) (float32, written out so GetExtN — a double log for float volumes —
//           runs once), per slice the column sums of the 32-row band cs(x, band, z), per segment the z sums zs(x,y,seg)
//           and their band sums;
//   mid    `sat_mid_kernel`: exclusive prefixes of those carries over bands / segments (tens of MB);
//   pass 2 `sat_scan_kernel`: per slice the tile-local 2D inclusive scan (x: 3 adds + a warp shuffle scan per row,
//           y: running sums), accumulated over z in registers (so the 32 accumulators of a thread hold the z-accumulated
//           tile-local scan); the carry from the bands above enters as an extra addend of the band's first row, the
//           carry from the slices before the segment as the accumulators' initial value; warps exchange their tile's
//           right edge and bottom row through shared memory (one barrier per slice); float32 store.
// All sums in double.  HBM traffic: 16 B/cell for float volumes (4 raw + 4 E, 4 E + 4 table), 10-11 B/cell for integer
// volumes, plus < 1 B/cell of carries — against 56 B/cell of the separable scan.  Needs w % 4 == 0 and w <= 512.
namespace sat2 {
constexpr int XW = 4, YW = 4, R = 8, C = 4;
constexpr int THREADS = 32 * XW * YW;       // 512
constexpr int XSPAN = 32 * XW * C;          // 512 columns per CTA
constexpr int BAND = YW * R;                // 32 rows per CTA
constexpr int STAGES = 2;

// log(v) for finite v >= 1, as accurate as the library's (measured error < 1.2 ulp of double) at a quarter of its FP64
// instructions: v = 2^e * m, the top 7 mantissa bits pick a centre c_k, r = m / c_k - 1 (one FMA with the tabulated
// 1/c_k, |r| <= 2^-7), log v = e ln2 + log c_k + log1p(r) with a degree-8 Taylor polynomial; the first bin has c = 1
// exactly, so values just above 1 — small opacities — lose nothing to cancellation.  GetExtN's (float)log(1/(1-a)) keeps
// its IEEE double division (for small a the reference's result is dominated by that rounding, so it is part of the
// parity contract); the float result equals the libm evaluation on all 3*10^7 opacities of a host-side sweep, and can
// differ only where the double value lies within ~1e-16 relative of a float rounding boundary.
struct LogTab { double invc, logc; };
constexpr int LOG_BINS = 128;
__device__ __forceinline__ double fast_log_ge1(double v, const LogTab* tab) {
  const int hi = __double2hiint(v);
  const int e = ((hi >> 20) & 0x7ff) - 1023;
  const int k = (hi >> 13) & (LOG_BINS - 1);
  const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(v));   // [1, 2)
  const LogTab t = tab[k];
  const double r = fma(m, t.invc, -1.0);
  double q = fma(r, -1.0 / 8.0, 1.0 / 7.0);
  q = fma(r, q, -1.0 / 6.0); q = fma(r, q, 1.0 / 5.0); q = fma(r, q, -1.0 / 4.0); q = fma(r, q, 1.0 / 3.0); q = fma(r, q, -0.5);
  const double p = fma(r * r, q, r);
  return fma((double)e, 0.69314718055994530942, t.logc) + p;
}
inline void build_log_table(LogTab* tab) {
  for (int k = 0; k < LOG_BINS; k++) {
    const long double c = 1.0L + ((long double)k + 0.5L) / (long double)LOG_BINS;
    double invc = (double)(1.0L / c);
    if (k == 0) invc = 1.0;                                  // c = 1: r = m - 1 exactly
    const long double c_used = 1.0L / (long double)invc;     // the centre the rounded reciprocal actually stands for
    tab[k].invc = invc; tab[k].logc = (double)logl(c_used);
  }
}
// TransferFunction1D::GetExtN of four normalised float samples (tf_get_extn above, same arithmetic) with fast_log_ge1 for
// the logarithm.  Written without data-dependent branches around the common path so the four evaluations interleave:
// the opacity of every sample first, then — only if one of them is non-zero — the four divisions and logarithms.
__device__ __forceinline__ void ext_of_float4(const double* __restrict__ table, int max_density, int ext_type, const float (&nv)[4],
                                              const LogTab* tab, float zero_below, float (&e)[4]) {
  // samples below the first transfer-function cell with a non-zero opacity: both lerp endpoints are 0, GetExtN = 0
  if (nv[0] < zero_below && nv[1] < zero_below && nv[2] < zero_below && nv[3] < zero_below) {
    e[0] = e[1] = e[2] = e[3] = 0.0f;
    return;
  }
  float a[4];
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const double value = (double)nv[c] * ((double)max_density / 1.0);
    const bool outside = value < 0.0f || value > (float)max_density;          // NaN: both false, as in tf_get_extn
    const bool at_max = fabs(value - (float)max_density) < 0.000001;
    const int iv = min(max((int)value, 0), max_density - 1);              // == (int)value whenever the lerp is selected
    const double t = value - (int)value;
    const double lo = table[4 * iv + 3], hi = table[4 * (iv + 1) + 3];
    const float lerp = (float)((1.0 - t) * lo + t * hi);
    a[c] = outside ? 0.0f : (at_max ? (float)hi : lerp);
  }
  if (ext_type || (a[0] == 0.0f && a[1] == 0.0f && a[2] == 0.0f && a[3] == 0.0f)) {
#pragma unroll
    for (int c = 0; c < 4; c++) e[c] = a[c];
    return;
  }
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const double v = 1.0 / (1.0 - (double)a[c]);
    const bool ok = a[c] > 0.0f && a[c] < 1.0f && v > 1.0;
    const double l = fast_log_ge1(ok ? v : 2.0, tab);
    e[c] = a[c] == 0.0f ? 0.0f : (float)l;
    if (!ok && a[c] != 0.0f) e[c] = (float)log(v);                             // inf / NaN / v == 1: as the library computes them
  }
}

template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem, const void* gmem, bool pred) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  const int n = pred ? BYTES : 0;                     // src-size 0: nothing is read, the destination is zero-filled
  if (BYTES == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(n) : "memory");
  else asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" ::"r"(s), "l"(gmem), "n"(BYTES), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <typename T> struct Raw4;
template <> struct alignas(4) Raw4<uint8_t> { uint8_t v[4]; };
template <> struct alignas(8) Raw4<uint16_t> { uint16_t v[4]; };
template <> struct alignas(16) Raw4<float> { float v[4]; };

struct Geo {
  int w, h, d, seglen, nbands, nseg;
};

// Elementwise E = GetExtN(sample) of a float volume, at full occupancy (small CTAs, 4 voxels per thread): the double lerp / division /
// logarithm of GetExtN is a long dependent chain, and inside the one-CTA-per-SM tile kernel below nothing hides it (ncu r1d: 0.69 of
// the 1.04 ms build, DRAM 19 % busy).  Here 8 CTAs per SM do.  HBM-bound: 4 B read + 4 B written per voxel.
__global__ void __launch_bounds__(256, 4) sat_ext_kernel(const float* __restrict__ raw, const double* __restrict__ table, int max_density, int ext_type,
                                                         size_t n4, const LogTab* __restrict__ logtab, float zero_below, float* __restrict__ E) {
  __shared__ LogTab s_log[LOG_BINS];
  for (int i = threadIdx.x; i < LOG_BINS; i += blockDim.x) s_log[i] = logtab[i];
  __syncthreads();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(raw) + i);
    const float nv[4] = {v.x, v.y, v.z, v.w};
    float e[4];
    ext_of_float4(table, max_density, ext_type, nv, s_log, zero_below, e);
    reinterpret_cast<float4*>(E)[i] = make_float4(e[0], e[1], e[2], e[3]);
  }
}

// MODE 0: extinction from the raw voxels (table for integers, GetExtN for floats), E written out;  MODE 1: `raw` already holds E
// (float), nothing written;  MODE 2: integer voxels through the table, nothing written (the scan pass looks the table up again).
template <typename T, int MODE>
__global__ void __launch_bounds__(THREADS, 1) sat_reduce_kernel(const T* __restrict__ raw, const float* __restrict__ lut,
                                                                const double* __restrict__ table, int max_density, int ext_type, Geo g,
                                                                float* __restrict__ E, double* __restrict__ cs, double* __restrict__ zs,
                                                                double* __restrict__ bandzs, const LogTab* __restrict__ logtab, float zero_below) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  using V = Raw4<T>;
  V* stage = reinterpret_cast<V*>(s_raw);                                                   // [STAGES][R][THREADS]
  double* red = reinterpret_cast<double*>(s_raw + (size_t)STAGES * R * THREADS * sizeof(V));  // [2][YW][XSPAN]
  LogTab* s_log = reinterpret_cast<LogTab*>(red + 2 * YW * XSPAN);                          // [LOG_BINS] (float volumes)
  if (sizeof(T) == 4 && MODE == 0) {
    for (int i = threadIdx.x; i < LOG_BINS; i += THREADS) s_log[i] = logtab[i];
    __syncthreads();
  }
  const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5, wx = wq & (XW - 1), wy = wq / XW;
  const int band = blockIdx.x, seg = blockIdx.y;
  const int xl = wx * 32 * C + lane * C, x0 = xl, yb = band * BAND + wy * R;
  const int z0 = seg * g.seglen, z1 = min(g.d, z0 + g.seglen);
  const bool xin = x0 < g.w;
  auto issue = [&](int z, int st) {
    if (z < z1) {
#pragma unroll
      for (int r = 0; r < R; r++) {
        const int y = yb + r;
        const bool p = xin && y < g.h;
        const T* src = p ? raw + ((size_t)x0 + (size_t)g.w * ((size_t)y + (size_t)g.h * (size_t)z)) : raw;
        cp_async<(int)sizeof(V)>(&stage[(st * R + r) * THREADS + tid], src, p);
      }
    }
    cp_commit();
  };
  double acc[R][C], bz[C];
#pragma unroll
  for (int r = 0; r < R; r++)
#pragma unroll
    for (int c = 0; c < C; c++) acc[r][c] = 0.0;
#pragma unroll
  for (int c = 0; c < C; c++) bz[c] = 0.0;
#pragma unroll
  for (int s = 0; s < STAGES; s++) issue(z0 + s, s);
  for (int z = z0; z < z1; z++) {
    const int st = (z - z0) % STAGES, par = (z - z0) & 1;
    cp_wait<STAGES - 1>();
    double col[C];
#pragma unroll
    for (int c = 0; c < C; c++) col[c] = 0.0;
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int y = yb + r;
      const bool p = xin && y < g.h;
      const V v = stage[(st * R + r) * THREADS + tid];
      float e[C];
      if (MODE == 1) {
#pragma unroll
        for (int c = 0; c < C; c++) e[c] = p ? (float)v.v[c] : 0.0f;
      } else if (sizeof(T) != 4) {
#pragma unroll
        for (int c = 0; c < C; c++) e[c] = p ? __ldg(lut + (unsigned)v.v[c]) : 0.0f;
      } else {
        const float nv4[C] = {(float)v.v[0], (float)v.v[1], (float)v.v[2], (float)v.v[3]};
        ext_of_float4(table, max_density, ext_type, nv4, s_log, zero_below, e);         // staged zeros outside the volume: computed, not stored
#pragma unroll
        for (int c = 0; c < C; c++) e[c] = p ? e[c] : 0.0f;
      }
      if (MODE == 0 && p) *reinterpret_cast<float4*>(E + ((size_t)x0 + (size_t)g.w * ((size_t)y + (size_t)g.h * (size_t)z))) = make_float4(e[0], e[1], e[2], e[3]);
#pragma unroll
      for (int c = 0; c < C; c++) { acc[r][c] += (double)e[c]; col[c] += (double)e[c]; }
    }
    double* rd = red + ((size_t)(par * YW + wy) * XSPAN + xl);
    *reinterpret_cast<double2*>(rd) = make_double2(col[0], col[1]);
    *reinterpret_cast<double2*>(rd + 2) = make_double2(col[2], col[3]);
    issue(z + STAGES, st);                      // refills the stage this thread has just consumed
    __syncthreads();
    if (wy == 0 && xin) {
      double s4[C] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int q = 0; q < YW; q++) {
        const double2 a = *reinterpret_cast<const double2*>(red + ((size_t)(par * YW + q) * XSPAN + xl));
        const double2 b = *reinterpret_cast<const double2*>(red + ((size_t)(par * YW + q) * XSPAN + xl + 2));
        s4[0] += a.x; s4[1] += a.y; s4[2] += b.x; s4[3] += b.y;
      }
      double* o = cs + ((size_t)z * g.nbands + band) * (size_t)g.w + x0;
      *reinterpret_cast<double2*>(o) = make_double2(s4[0], s4[1]);
      *reinterpret_cast<double2*>(o + 2) = make_double2(s4[2], s4[3]);
#pragma unroll
      for (int c = 0; c < C; c++) bz[c] += s4[c];
    }
  }
  cp_wait<0>();
  if (xin) {
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int y = yb + r;
      if (y < g.h) {
        double* o = zs + ((size_t)seg * g.h + y) * (size_t)g.w + x0;
        *reinterpret_cast<double2*>(o) = make_double2(acc[r][0], acc[r][1]);
        *reinterpret_cast<double2*>(o + 2) = make_double2(acc[r][2], acc[r][3]);
      }
    }
    if (wy == 0) {
      double* o = bandzs + ((size_t)seg * g.nbands + band) * (size_t)g.w + x0;
      *reinterpret_cast<double2*>(o) = make_double2(bz[0], bz[1]);
      *reinterpret_cast<double2*>(o + 2) = make_double2(bz[2], bz[3]);
    }
  }
}

// Exclusive prefix of n values `stride` apart, in place; eight loads in flight at a time.
__device__ __forceinline__ void prefix_in_place(double* p, int n, size_t stride) {
  double run = 0.0;
  for (int b0 = 0; b0 < n; b0 += 8) {
    double t[8];
#pragma unroll
    for (int j = 0; j < 8; j++) t[j] = b0 + j < n ? p[(size_t)(b0 + j) * stride] : 0.0;
#pragma unroll
    for (int j = 0; j < 8; j++) { if (b0 + j < n) p[(size_t)(b0 + j) * stride] = run; run += t[j]; }
  }
}

// Exclusive prefixes of the carries: cs over bands (per x, z) and zs over segments (per x, y) in place, bandzs over
// bands and segments into zsy.
__global__ void __launch_bounds__(256) sat_mid_kernel(double* __restrict__ cs, double* __restrict__ zs, const double* __restrict__ bandzs,
                                                      double* __restrict__ zsy, Geo g, unsigned blocks_a, unsigned blocks_b) {
  const size_t w = (size_t)g.w;
  if (blockIdx.x < blocks_a) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= w * g.d) return;
    const size_t x = i % w, z = i / w;
    prefix_in_place(cs + z * g.nbands * w + x, g.nbands, w);
  } else if (blockIdx.x < blocks_a + blocks_b) {
    const size_t i = (size_t)(blockIdx.x - blocks_a) * 256 + threadIdx.x;
    if (i >= w * g.h) return;
    prefix_in_place(zs + i, g.nseg, (size_t)g.h * w);
  } else {
    // zsy(x, band, seg) = sum of bandzs over bands < band and segments < seg; read-only input, so the loads pipeline
    const size_t i = (size_t)(blockIdx.x - blocks_a - blocks_b) * 256 + threadIdx.x;
    if (i >= w * g.nbands) return;
    const size_t x = i % w; const int band = (int)(i / w);
    double run = 0.0;
    for (int s = 0; s < g.nseg; s++) {
      zsy[((size_t)s * g.nbands + band) * w + x] = run;
      double t = 0.0;
      for (int b = 0; b < band; b++) t += bandzs[((size_t)s * g.nbands + b) * w + x];
      run += t;
    }
  }
}

__device__ __forceinline__ double warp_incl_scan(double v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
  return v;
}

// One row of the tile-local 2D scan: a[] = the row's 4 values; x-scan over the warp's 128 columns, then onto the
// running column sums run[] (y-scan), then onto the z accumulators.
__device__ __forceinline__ void scan_row(double (&a)[C], double (&run)[C], double (&acc)[C], int lane) {
  a[1] += a[0]; a[2] += a[1]; a[3] += a[2];
  const double incl = warp_incl_scan(a[3], lane);
  double excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = 0.0;
#pragma unroll
  for (int c = 0; c < C; c++) { run[c] += a[c] + excl; acc[c] += run[c]; }
}

// T = float: E holds the extinctions.  T = uint8_t / uint16_t: E holds the raw voxels and the extinction is looked up again in `lut`
// (256 entries in shared memory, 65536 through the read-only cache) — the float copy of the volume is never written.
template <typename T>
__global__ void __launch_bounds__(THREADS, 1) sat_scan_kernel(const T* __restrict__ E, const float* __restrict__ lut, const double* __restrict__ csx, const double* __restrict__ ZS,
                                                              const double* __restrict__ ZSY, Geo g, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  using V = Raw4<T>;
  V* stage = reinterpret_cast<V*>(s_raw);                                             // [STAGES][R][THREADS]
  double2* cstage = reinterpret_cast<double2*>(s_raw + (size_t)STAGES * R * THREADS * 16);   // [STAGES][2][32 * XW] (wy == 0 warps)
  double* right = reinterpret_cast<double*>(cstage + STAGES * 2 * 32 * XW);          // [2][YW][R][XW]
  double* bottom = right + 2 * YW * R * XW;                                           // [2][YW][XSPAN]
  float* s_lut = reinterpret_cast<float*>(bottom + 2 * YW * XSPAN);                   // [256] (8-bit volumes)
  if (sizeof(T) == 1) {
    for (int i = threadIdx.x; i < 256; i += THREADS) s_lut[i] = lut[i];
    __syncthreads();
  }
  const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5, wx = wq & (XW - 1), wy = wq / XW;
  const int band = blockIdx.x, seg = blockIdx.y;
  const int xl = wx * 32 * C + lane * C, x0 = xl, yb = band * BAND + wy * R;
  const int z0 = seg * g.seglen, z1 = min(g.d, z0 + g.seglen);
  const bool xin = x0 < g.w;
  const int W = g.w + 2, H = g.h + 2;
  auto issue = [&](int z, int st) {
    if (z < z1) {
#pragma unroll
      for (int r = 0; r < R; r++) {
        const int y = yb + r;
        const bool p = xin && y < g.h;
        const T* src = p ? E + ((size_t)x0 + (size_t)g.w * ((size_t)y + (size_t)g.h * (size_t)z)) : E;
        cp_async<(int)sizeof(V)>(&stage[(st * R + r) * THREADS + tid], src, p);
      }
      if (wy == 0) {
        const double* src = xin ? csx + ((size_t)z * g.nbands + band) * (size_t)g.w + x0 : csx;
        cp_async<16>(&cstage[(st * 2 + 0) * 32 * XW + wx * 32 + lane], src, xin);
        cp_async<16>(&cstage[(st * 2 + 1) * 32 * XW + wx * 32 + lane], xin ? src + 2 : src, xin);
      }
    }
    cp_commit();
  };
#pragma unroll
  for (int s = 0; s < STAGES; s++) issue(z0 + s, s);
  // accumulators start at the tile-local scan of the z carry (sum of the slices before the segment), its first row
  // carrying the band carry of those slices
  double acc[R][C];
  {
    double run[C] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int y = yb + r;
      double a[C] = {0.0, 0.0, 0.0, 0.0};
      if (xin && y < g.h) {
        const double* p = ZS + ((size_t)seg * g.h + y) * (size_t)g.w + x0;
        const double2 u = *reinterpret_cast<const double2*>(p), v = *reinterpret_cast<const double2*>(p + 2);
        a[0] = u.x; a[1] = u.y; a[2] = v.x; a[3] = v.y;
      }
      if (r == 0 && wy == 0 && xin) {
        const double* p = ZSY + ((size_t)seg * g.nbands + band) * (size_t)g.w + x0;
        const double2 u = *reinterpret_cast<const double2*>(p), v = *reinterpret_cast<const double2*>(p + 2);
        a[0] += u.x; a[1] += u.y; a[2] += v.x; a[3] += v.y;
      }
#pragma unroll
      for (int c = 0; c < C; c++) acc[r][c] = 0.0;
      scan_row(a, run, acc[r], lane);
    }
  }
  for (int z = z0; z < z1; z++) {
    const int st = (z - z0) % STAGES, par = (z - z0) & 1;
    cp_wait<STAGES - 1>();
    {
      double run[C] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int r = 0; r < R; r++) {
        const V v = stage[(st * R + r) * THREADS + tid];            // zero-filled outside the volume
        double a[C];
        if (sizeof(T) == 4) {
#pragma unroll
          for (int c = 0; c < C; c++) a[c] = (double)v.v[c];
        } else {
          const bool p = xin && yb + r < g.h;                       // a zero-filled voxel is voxel value 0, whose extinction need not be 0
#pragma unroll
          for (int c = 0; c < C; c++) a[c] = p ? (double)(sizeof(T) == 1 ? s_lut[(unsigned)v.v[c]] : __ldg(lut + (unsigned)v.v[c])) : 0.0;
        }
        if (r == 0 && wy == 0) {
          const double2 u = cstage[(st * 2 + 0) * 32 * XW + wx * 32 + lane], t = cstage[(st * 2 + 1) * 32 * XW + wx * 32 + lane];
          a[0] += u.x; a[1] += u.y; a[2] += t.x; a[3] += t.y;
        }
        scan_row(a, run, acc[r], lane);
      }
    }
    // tile edges of the z-accumulated local scan
    if (lane == 31) {
#pragma unroll
      for (int r = 0; r < R; r++) right[((par * YW + wy) * R + r) * XW + wx] = acc[r][C - 1];
    }
    {
      double* b = bottom + (size_t)(par * YW + wy) * XSPAN + xl;
      *reinterpret_cast<double2*>(b) = make_double2(acc[R - 1][0], acc[R - 1][1]);
      *reinterpret_cast<double2*>(b + 2) = make_double2(acc[R - 1][2], acc[R - 1][3]);
    }
    issue(z + STAGES, st);
    __syncthreads();
    double corner = 0.0, colc[C] = {0.0, 0.0, 0.0, 0.0};
    for (int q = 0; q < wy; q++) {
      for (int k = 0; k < wx; k++) corner += right[((par * YW + q) * R + (R - 1)) * XW + k];
      const double* b = bottom + (size_t)(par * YW + q) * XSPAN + xl;
      const double2 u = *reinterpret_cast<const double2*>(b), v = *reinterpret_cast<const double2*>(b + 2);
      colc[0] += u.x; colc[1] += u.y; colc[2] += v.x; colc[3] += v.y;
    }
#pragma unroll
    for (int c = 0; c < C; c++) colc[c] += corner;
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int y = yb + r;
      double rowc = 0.0;
      for (int k = 0; k < wx; k++) rowc += right[((par * YW + wy) * R + r) * XW + k];
      if (xin && y < g.h) {
        float* o = out + ((size_t)(x0 + 1) + (size_t)W * ((size_t)(y + 1) + (size_t)H * (size_t)(z + 1)));
#pragma unroll
        for (int c = 0; c < C; c++) o[c] = (float)(acc[r][c] + (rowc + colc[c]));
      }
    }
  }
  cp_wait<0>();
}

// The zero planes X/Y/Z = 0 and the repeated planes X = w+1, Y = h+1, Z = d+1 of the bordered table.
__global__ void __launch_bounds__(256) sat_border_kernel(float* __restrict__ out, int w, int h, int d) {
  const int W = w + 2, H = h + 2, D = d + 2;
  const size_t nx = (size_t)H * D, ny = (size_t)W * D, nz = (size_t)W * H;
  size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  int X, Y, Z;
  if (i < 2 * nx) { X = i < nx ? 0 : W - 1; i %= nx; Y = (int)(i % H); Z = (int)(i / H); }
  else if (i < 2 * nx + 2 * ny) { i -= 2 * nx; Y = i < ny ? 0 : H - 1; i %= ny; X = (int)(i % W); Z = (int)(i / W); }
  else if (i < 2 * nx + 2 * ny + 2 * nz) { i -= 2 * nx + 2 * ny; Z = i < nz ? 0 : D - 1; i %= nz; X = (int)(i % W); Y = (int)(i / W); }
  else return;
  float v = 0.0f;
  if (X > 0 && Y > 0 && Z > 0) v = out[(size_t)min(X, w) + (size_t)W * ((size_t)min(Y, h) + (size_t)H * (size_t)min(Z, d))];
  out[(size_t)X + (size_t)W * ((size_t)Y + (size_t)H * (size_t)Z)] = v;
}
}  // namespace sat2

// Footprint texture of the float32 table: texel (x,y,z) = {S(x,y), S(x+1,y), S(x,y+1), S(x+1,y+1)} of slice z,
// edge-clamped; the shader's GL_LINEAR fetch of the R32F SAT (ebsrenderer.cpp:716) becomes two point fetches and a
// float32 lerp in the oracle's order (the hardware's 1.8 fixed-point weights are useless on values of this size).
// A thread walks QUAD_ROWS rows of one column: two loads per row (the lower pair of a texel is the upper pair of the next), all issued
// before the first store — 2.2 GB of surface stores at 514^3 (1.04 ms, 2.1 TB/s of stores; one row per thread: 1.15 ms).
constexpr int QUAD_ROWS = 8;
__global__ void __launch_bounds__(128) sat_quad_kernel(const float* __restrict__ s, cudaSurfaceObject_t dst, int W, int H, int D) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y0 = blockIdx.y * QUAD_ROWS, z = blockIdx.z;
  if (x >= W) return;
  const int x1 = min(x + 1, W - 1);
  const float* p = s + (size_t)W * H * z;
  float a[QUAD_ROWS + 1], b[QUAD_ROWS + 1];
#pragma unroll
  for (int r = 0; r <= QUAD_ROWS; r++) {
    const int y = min(y0 + r, H - 1);
    a[r] = __ldg(p + (size_t)W * y + x); b[r] = __ldg(p + (size_t)W * y + x1);
  }
#pragma unroll
  for (int r = 0; r < QUAD_ROWS; r++)
    if (y0 + r < H) surf3Dwrite(make_float4(a[r], b[r], a[r + 1], b[r + 1]), dst, x * 16, y0 + r, z);
}

void ExtSat::release() {
  if (tex_quad) cudaDestroyTextureObject(tex_quad);
  if (quad) cudaFreeArray(quad);
  if (sat) cudaFree(sat);
  tex_quad = 0; quad = nullptr; sat = nullptr; W = H = D = 0; vol_gen = tf_gen = 0;
}

int ExtSat::ensure(vrb_ctx* ctx, int fast) {
  const Volume& v = ctx->vol;
  const TransferFunction& t = ctx->tf;
  if (sat && vol_gen == v.generation && tf_gen == t.generation && built_fast == fast) return VRB_OK;
  release();
  VRB_REQUIRE(v.raw && t.table, VRB_ERR_INVALID, "extinction SAT: volume and transfer function must be uploaded first");
  W = (int)v.w + 2; H = (int)v.h + 2; D = (int)v.d + 2;
  const size_t n = (size_t)W * H * D;
  cudaStream_t s = ctx->stream;
  double* acc = nullptr; float* lut = nullptr;
  float* E = nullptr; double *cs = nullptr, *zs = nullptr, *bzs = nullptr; sat2::LogTab* logtab = nullptr; float zero_below = -INFINITY;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  auto body = [&]() -> int {
    const bool v2 = fast == 1 && (v.w % 4) == 0 && v.w <= sat2::XSPAN;
    if (!v2) VRB_CUDA(cudaMalloc(&acc, n * sizeof(double)));
    VRB_CUDA(cudaMalloc(&sat, n * sizeof(float)));
    VRB_CUDA(cudaEventCreate(&e0)); VRB_CUDA(cudaEventCreate(&e1));
    const int maxd = t.n - 1;
    if (v.dtype != VRB_F32) {
      // GetExtN of every representable voxel value, with the host's libm
      std::vector<double> table((size_t)t.n * 4);
      VRB_CUDA(cudaMemcpyAsync(table.data(), t.table, table.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
      VRB_CUDA(cudaStreamSynchronize(s));
      const int nv = v.dtype == VRB_U8 ? 256 : 65536;
      const double maxv = v.dtype == VRB_U8 ? 256.0 - 1.0 : 65536.0 - 1.0;
      std::vector<float> h((size_t)nv);
      for (int i = 0; i < nv; i++) h[i] = tf_get_extn(table.data(), maxd, t.extinction_type, (double)i / maxv);
      VRB_CUDA(cudaMalloc(&lut, h.size() * sizeof(float)));
      VRB_CUDA(cudaMemcpyAsync(lut, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice, s));
      VRB_CUDA(cudaStreamSynchronize(s));
    }
    sat2::Geo g{};
    if (v2) {
      g.w = (int)v.w; g.h = (int)v.h; g.d = (int)v.d;
      g.nbands = (g.h + sat2::BAND - 1) / sat2::BAND;
      // one CTA per SM (512 threads, ~170 KB shared memory): bands x segments = one wave, segments of >= 8 slices
      g.nseg = std::max(1, std::min(std::max(1, ctx->sm_count / g.nbands), std::max(1, g.d / 8)));
      g.seglen = (g.d + g.nseg - 1) / g.nseg;
      g.nseg = (g.d + g.seglen - 1) / g.seglen;
      const size_t wz = (size_t)g.w;
      if (v.dtype == VRB_F32) VRB_CUDA(cudaMalloc(&E, wz * g.h * g.d * sizeof(float)));    // integer volumes: the scan pass looks the table up again
      VRB_CUDA(cudaMalloc(&cs, wz * g.nbands * g.d * sizeof(double)));
      VRB_CUDA(cudaMalloc(&zs, wz * g.h * g.nseg * sizeof(double)));
      VRB_CUDA(cudaMalloc(&bzs, 2 * wz * g.nbands * g.nseg * sizeof(double)));
      if (v.dtype == VRB_F32) {
        // largest normalised sample bound below which GetExtN is exactly 0: value = nv * maxd < j, table[0..j].a == 0
        std::vector<double> tb((size_t)t.n * 4);
        VRB_CUDA(cudaMemcpyAsync(tb.data(), t.table, tb.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
        VRB_CUDA(cudaStreamSynchronize(s));
        int j = -1;
        while (j + 1 < t.n && tb[4 * (size_t)(j + 1) + 3] == 0.0) j++;
        if (j >= 1) {
          zero_below = (float)((double)j / (double)maxd);
          while ((double)zero_below * (double)maxd > (double)j) zero_below = std::nextafterf(zero_below, -1.0f);
        }
        sat2::LogTab h[sat2::LOG_BINS];
        sat2::build_log_table(h);
        VRB_CUDA(cudaMalloc(&logtab, sizeof(h)));
        VRB_CUDA(cudaMemcpyAsync(logtab, h, sizeof(h), cudaMemcpyHostToDevice, s));
        VRB_CUDA(cudaStreamSynchronize(s));
      }
    }
    VRB_CUDA(cudaEventRecord(e0, s));
    if (v2) {
      const size_t wz = (size_t)g.w;
      double* zsy = bzs + wz * g.nbands * g.nseg;
      const dim3 grid(g.nbands, g.nseg);
      const size_t esz = v.dtype == VRB_U8 ? 4 : (v.dtype == VRB_U16 ? 8 : 16);
      const size_t sm1 = (size_t)sat2::STAGES * sat2::R * sat2::THREADS * esz + (size_t)2 * sat2::YW * sat2::XSPAN * sizeof(double) +
                         sat2::LOG_BINS * sizeof(sat2::LogTab);
      const size_t sm2 = (size_t)sat2::STAGES * sat2::R * sat2::THREADS * 16 + (size_t)sat2::STAGES * 2 * 32 * sat2::XW * 16 +
                         (size_t)2 * sat2::YW * sat2::R * sat2::XW * sizeof(double) + (size_t)2 * sat2::YW * sat2::XSPAN * sizeof(double) + 256 * sizeof(float);
      const int maxd2 = t.n - 1;
      const unsigned ba = (unsigned)((wz * g.d + 255) / 256), bb = (unsigned)((wz * g.h + 255) / 256), bc = (unsigned)((wz * g.nbands + 255) / 256);
      if (v.dtype == VRB_U8) {
        VRB_CUDA(cudaFuncSetAttribute(sat2::sat_reduce_kernel<uint8_t, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
        sat2::sat_reduce_kernel<uint8_t, 2><<<grid, sat2::THREADS, sm1, s>>>((const uint8_t*)v.raw, lut, t.table, maxd2, t.extinction_type, g, nullptr, cs, zs, bzs, logtab, zero_below);
      } else if (v.dtype == VRB_U16) {
        VRB_CUDA(cudaFuncSetAttribute(sat2::sat_reduce_kernel<uint16_t, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
        sat2::sat_reduce_kernel<uint16_t, 2><<<grid, sat2::THREADS, sm1, s>>>((const uint16_t*)v.raw, lut, t.table, maxd2, t.extinction_type, g, nullptr, cs, zs, bzs, logtab, zero_below);
      } else {
        // float volumes: GetExtN elementwise at full occupancy, then the carries from E
        const size_t n4 = wz * g.h * g.d / 4;                 // w % 4 == 0
        sat2::sat_ext_kernel<<<ctx->sm_count * 8, 256, 0, s>>>((const float*)v.raw, t.table, maxd2, t.extinction_type, n4, logtab, zero_below, E);
        VRB_CUDA(cudaFuncSetAttribute(sat2::sat_reduce_kernel<float, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
        sat2::sat_reduce_kernel<float, 1><<<grid, sat2::THREADS, sm1, s>>>(E, nullptr, t.table, maxd2, t.extinction_type, g, nullptr, cs, zs, bzs, logtab, zero_below);
      }
      VRB_CUDA(cudaGetLastError());
      sat2::sat_mid_kernel<<<ba + bb + bc, 256, 0, s>>>(cs, zs, bzs, zsy, g, ba, bb);
      if (v.dtype == VRB_U8) {
        VRB_CUDA(cudaFuncSetAttribute(sat2::sat_scan_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
        sat2::sat_scan_kernel<uint8_t><<<grid, sat2::THREADS, sm2, s>>>((const uint8_t*)v.raw, lut, cs, zs, zsy, g, sat);
      } else if (v.dtype == VRB_U16) {
        VRB_CUDA(cudaFuncSetAttribute(sat2::sat_scan_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
        sat2::sat_scan_kernel<uint16_t><<<grid, sat2::THREADS, sm2, s>>>((const uint16_t*)v.raw, lut, cs, zs, zsy, g, sat);
      } else {
        VRB_CUDA(cudaFuncSetAttribute(sat2::sat_scan_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
        sat2::sat_scan_kernel<float><<<grid, sat2::THREADS, sm2, s>>>(E, nullptr, cs, zs, zsy, g, sat);
      }
      const size_t faces = 2 * ((size_t)H * D + (size_t)W * D + (size_t)W * H);
      sat2::sat_border_kernel<<<(unsigned)((faces + 255) / 256), 256, 0, s>>>(sat, g.w, g.h, g.d);
    } else {
    dim3 fb(256), fg((W + 255) / 256, H, D);
    if (v.dtype == VRB_U8) sat_fill_kernel<uint8_t><<<fg, fb, 0, s>>>((const uint8_t*)v.raw, lut, t.table, maxd, t.extinction_type, (int)v.w, (int)v.h, (int)v.d, acc);
    else if (v.dtype == VRB_U16) sat_fill_kernel<uint16_t><<<fg, fb, 0, s>>>((const uint16_t*)v.raw, lut, t.table, maxd, t.extinction_type, (int)v.w, (int)v.h, (int)v.d, acc);
    else sat_fill_kernel<float><<<fg, fb, 0, s>>>((const float*)v.raw, nullptr, t.table, maxd, t.extinction_type, (int)v.w, (int)v.h, (int)v.d, acc);
    VRB_CUDA(cudaGetLastError());
    if (fast) {
      const size_t rows = (size_t)H * D;
      sat_scan_x_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(acc, W, rows);
      sat_scan_y_kernel<<<dim3((W + 255) / 256, D), 256, 0, s>>>(acc, W, H, D);
      sat_scan_z_kernel<<<dim3((W + 255) / 256, H), 256, 0, s>>>(acc, sat, W, H, D);
    } else {
      const int ntx = (W + WT_X - 1) / WT_X, nty = (H + WT_Y - 1) / WT_Y, ntz = (D + WT_Z - 1) / WT_Z;
      const size_t wsm = (size_t)(WT_Z + 1) * WT_SY * WT_SX * sizeof(double);
      VRB_CUDA(cudaFuncSetAttribute(sat_tile_wavefront_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm));
      for (int kd = 0; kd <= ntx + nty + ntz - 3; kd++)
        sat_tile_wavefront_kernel<<<ntx * nty, 256, wsm, s>>>(acc, W, H, D, kd, ntx, ntz);
      sat_to_float_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(acc, sat, n);
    }
    }
    VRB_CUDA(cudaGetLastError());
    VRB_CUDA(cudaEventRecord(e1, s));
    // sampling structure
    cudaChannelFormatDesc desc = cudaCreateChannelDesc<float4>();
    VRB_CUDA(cudaMalloc3DArray(&quad, &desc, make_cudaExtent((size_t)W, (size_t)H, (size_t)D), cudaArraySurfaceLoadStore));
    cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = quad;
    cudaSurfaceObject_t dst = 0;
    VRB_CUDA(cudaCreateSurfaceObject(&dst, &rd));
    cudaTextureDesc td{};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    cudaError_t e = cudaCreateTextureObject(&tex_quad, &rd, &td, nullptr);
    if (e == cudaSuccess) {
      sat_quad_kernel<<<dim3((W + 127) / 128, (H + QUAD_ROWS - 1) / QUAD_ROWS, D), 128, 0, s>>>(sat, dst, W, H, D);
      e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    }
    cudaDestroySurfaceObject(dst);
    VRB_CUDA(e);
    float ms = 0.f;
    VRB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    build_ms = ms;
    return VRB_OK;
  };
  const int rc = body();
  if (acc) cudaFree(acc);
  if (E) cudaFree(E);
  if (cs) cudaFree(cs);
  if (zs) cudaFree(zs);
  if (bzs) cudaFree(bzs);
  if (logtab) cudaFree(logtab);
  if (lut) cudaFree(lut);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  if (rc != VRB_OK) { release(); return rc; }
  vol_gen = v.generation; tf_gen = t.generation; built_fast = fast;
  return VRB_OK;
}

}  // namespace vrb
