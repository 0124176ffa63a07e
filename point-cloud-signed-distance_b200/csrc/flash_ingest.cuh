// flash_ingest.cuh -- point ingestion kernels of libflashsdf (row F3 of SURVEY 8f; the reference's side of this is
// `read_point_cloud` src/depthdata.jl:19-30 and the `sensed_points` a CostFunctor captures, src/gradientdescent.jl:41-47).
//
// A frame arrives as [n][3] fp64 or fp32 AoS (the memory of a Julia Vector{SVector{3,Float64}}).  The cost is a sum
// over points, so their order is free, and the cost kernel's unit of work is a GROUP of 32 consecutive points whose
// bounding ball drives every group-level cull and certificate (flash_dev.h group_prepass).  The tighter the balls, the
// more the kernel can decide once per group instead of once per point.  Pipeline, all on the ctx's copy stream:
//
//   frames_bounds_kernel   AABB of each frame (ordered-integer atomics)
//   points_keys_kernel     key = frame << hb*3 | Hilbert index of the point's cell in a 2^hb-per-axis CUBIC grid over the
//                          frame's AABB (hb = 10 bits per axis for up to 4 frames)
//   radix_*_kernel         stable LSD radix sort of (key, index), 8 bits per pass, 2-4 passes: per-warp digit histograms,
//                          one exclusive scan of the [digit][warp] table, stable scatter ranked with __match_any_sync
//   points_gather_kernel   one warp per group: float4 points in sorted order (each frame padded to whole groups), the
//                          permutation back to the caller's index (argmin output), and the group's bounding ball by
//                          Badoiu-Clarkson iterations on the warp (lanes = points)
//
// Measured on K5 (10 M uniform points, 32-point groups): the round-1 order (one counting sort into 64^3 Morton cells,
// arbitrary order inside a cell, ball = half diagonal of the bounding box) gives a median ball radius of 2.49 cm; the
// Hilbert order with Badoiu-Clarkson balls 1.64 cm (a k-d tree with exact 32-point leaves: 1.43 cm).  The stable sort also
// makes the point order -- and with it the summation order of the deterministic mode -- reproducible.
#pragma once
#include <cuda_runtime.h>

#include "flash_kernels.cuh"

namespace flashk {

constexpr int MAX_FRAMES = 64;

// per-frame layout of a point set: frame f holds cnt[f] points at [off[f], off[f] + cnt[f]) of the sorted buffers,
// off[f] a multiple of 32 (groups never straddle frames); src_off[f] is its start in the caller-order staging buffer
struct FrameTable {
  int n_frames;
  long long off[MAX_FRAMES + 1];
  long long cnt[MAX_FRAMES];
  long long src_off[MAX_FRAMES + 1];
};

// Skilling's transpose form of the Hilbert index, `bits` per axis (<= 10), returned interleaved (x highest)
__device__ __forceinline__ unsigned hilbert_index(unsigned x, unsigned y, unsigned z, int bits) {
  unsigned X[3] = {x, y, z};
  const unsigned M = 1u << (bits - 1);
  for (unsigned Q = M; Q > 1u; Q >>= 1) {
    const unsigned P = Q - 1u;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      if (X[i] & Q) X[0] ^= P;
      else { const unsigned t = (X[0] ^ X[i]) & P; X[0] ^= t; X[i] ^= t; }
    }
  }
  X[1] ^= X[0]; X[2] ^= X[1];
  unsigned t = 0u;
  for (unsigned Q = M; Q > 1u; Q >>= 1) if (X[2] & Q) t ^= Q - 1u;
  X[0] ^= t; X[1] ^= t; X[2] ^= t;
  unsigned key = 0u;
  for (int b = bits - 1; b >= 0; --b) key = (key << 3) | (((X[0] >> b) & 1u) << 2) | (((X[1] >> b) & 1u) << 1) | ((X[2] >> b) & 1u);
  return key;
}

// bounds[f][0..2] = min, [3..5] = max (ordered bits); blockIdx.y = frame
template <class T>
__global__ void frames_bounds_kernel(const T* __restrict__ src, const FrameTable ft, unsigned* __restrict__ bounds) {
  const int f = blockIdx.y;
  const T* s = src + 3 * ft.src_off[f];
  const long long n = ft.cnt[f];
  float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float3 p = load_xyz(s, i);
    if (p.x == p.x && p.y == p.y && p.z == p.z && fabsf(p.x) < 1e30f && fabsf(p.y) < 1e30f && fabsf(p.z) < 1e30f) {
      lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
      hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o)); hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o)); }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { atomicMin(bounds + 6 * f + k, ordered_bits(lo[k])); atomicMax(bounds + 6 * f + 3 + k, ordered_bits(hi[k])); }
  }
}
__global__ void bounds_init_kernel(unsigned* __restrict__ bounds, int n_frames, unsigned* __restrict__ scan_done) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 6 * n_frames) bounds[i] = (i % 6) < 3 ? 0xffffffffu : 0u;
  if (i == 0) *scan_done = 0u;
}

// key of every point: (frame << 3*hb | Hilbert index) >> drop, and its index in the staging buffer
template <class T>
__global__ void points_keys_kernel(const T* __restrict__ src, const FrameTable ft, const unsigned* __restrict__ bounds, int hb, int drop,
                                   unsigned* __restrict__ keys, int* __restrict__ idx) {
  const int f = blockIdx.y;
  const long long n = ft.cnt[f];
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long g = ft.src_off[f] + i;
  const float3 p = load_xyz(src, g);
  const unsigned* b = bounds + 6 * f;
  const float lo[3] = {from_ordered_bits(b[0]), from_ordered_bits(b[1]), from_ordered_bits(b[2])};
  const float ext = fmaxf(fmaxf(from_ordered_bits(b[3]) - lo[0], from_ordered_bits(b[4]) - lo[1]), from_ordered_bits(b[5]) - lo[2]);
  const float sc = ext > 0.f ? (float)(1 << hb) / ext : 0.f;
  const float v[3] = {p.x, p.y, p.z};
  unsigned c[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    int q = (int)((v[k] - lo[k]) * sc);          // NaN -> 0 after the clamp
    q = q < 0 ? 0 : (q > (1 << hb) - 1 ? (1 << hb) - 1 : q);
    c[k] = (unsigned)q;
  }
  const unsigned key = ((unsigned)f << (3 * hb)) | hilbert_index(c[0], c[1], c[2], hb);
  keys[g] = key >> drop;
  idx[g] = (int)g;
}

// ---- stable LSD radix sort, 8 bits per pass.  Every WARP owns a contiguous chunk of E keys (E a multiple of 32).
constexpr int RS_BLOCK = 256;
constexpr int RS_WARPS = RS_BLOCK / 32;
__global__ void __launch_bounds__(RS_BLOCK) radix_hist_kernel(const unsigned* __restrict__ keys, long long n, int shift, int E, int nW, int* __restrict__ table) {
  __shared__ int cnt[RS_WARPS][256];
  const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
  const long long w = (long long)blockIdx.x * RS_WARPS + wl;
  for (int d = lane; d < 256; d += 32) cnt[wl][d] = 0;
  __syncwarp();
  if (w < nW) {
    const long long b0 = w * E;
    long long b1 = b0 + E; if (b1 > n) b1 = n;
    for (long long base = b0; base < b1; base += 32) {
      const long long i = base + lane;
      const bool on = i < b1;
      const unsigned d = on ? (keys[i] >> shift) & 255u : 256u + lane;   // inactive lanes get unique pseudo-digits
      const unsigned peers = __match_any_sync(0xffffffffu, d);
      if (on && (__ffs(peers) - 1) == lane) cnt[wl][d] += __popc(peers);
      __syncwarp();
    }
    for (int d = lane; d < 256; d += 32) table[(long long)d * nW + w] = cnt[wl][d];
  }
}
// exclusive scan of `table` (len elements): every block scans 1024 elements locally; the LAST block to finish scans the
// block sums (exclusive, in place).  A scattered element i then starts at table[i] + sums[i >> 10].
__global__ void __launch_bounds__(1024) radix_scan_kernel(int* __restrict__ table, long long len, int* __restrict__ sums, unsigned* __restrict__ done) {
  __shared__ int wsum[32];
  __shared__ bool last;
  const long long i = (long long)blockIdx.x * 1024 + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int v = i < len ? table[i] : 0;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) wsum[w] = x;
  __syncthreads();
  if (w == 0) {
    int s = wsum[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
    wsum[lane] = s;
  }
  __syncthreads();
  const int incl = x + (w > 0 ? wsum[w - 1] : 0);
  if (i < len) table[i] = incl - v;
  if (threadIdx.x == 1023) {
    sums[blockIdx.x] = incl;
    __threadfence();
    last = atomicAdd(done, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  // this block scans the gridDim.x block sums
  int carry = 0;
  for (int base = 0; base < (int)gridDim.x; base += 1024) {
    const int j = base + threadIdx.x;
    const int sv = j < (int)gridDim.x ? ((volatile int*)sums)[j] : 0;
    int sx = sv;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, sx, o); if (lane >= o) sx += y; }
    __syncthreads();
    if (lane == 31) wsum[w] = sx;
    __syncthreads();
    if (w == 0) {
      int s = wsum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
      wsum[lane] = s;
    }
    __syncthreads();
    const int inc2 = sx + (w > 0 ? wsum[w - 1] : 0) + carry;
    if (j < (int)gridDim.x) sums[j] = inc2 - sv;
    carry += wsum[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) *done = 0u;   // ready for the next pass
}
__global__ void __launch_bounds__(RS_BLOCK) radix_scatter_kernel(const unsigned* __restrict__ keys_in, const int* __restrict__ idx_in, long long n, int shift, int E,
                                                                 int nW, const int* __restrict__ table, const int* __restrict__ sums,
                                                                 unsigned* __restrict__ keys_out, int* __restrict__ idx_out) {
  __shared__ int pos[RS_WARPS][256];
  const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
  const long long w = (long long)blockIdx.x * RS_WARPS + wl;
  if (w >= nW) return;
  for (int d = lane; d < 256; d += 32) { const long long t = (long long)d * nW + w; pos[wl][d] = table[t] + sums[t >> 10]; }
  __syncwarp();
  const long long b0 = w * E;
  long long b1 = b0 + E; if (b1 > n) b1 = n;
  const unsigned lt = (1u << lane) - 1u;
  for (long long base = b0; base < b1; base += 32) {
    const long long i = base + lane;
    const bool on = i < b1;
    const unsigned k = on ? keys_in[i] : 0u;
    const unsigned d = on ? (k >> shift) & 255u : 256u + lane;
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    int at = 0;
    if (on) at = pos[wl][d] + __popc(peers & lt);      // stable: lower lanes = earlier elements first
    __syncwarp();
    if (on && (__ffs(peers) - 1) == lane) pos[wl][d] += __popc(peers);
    __syncwarp();
    if (on) { keys_out[at] = k; idx_out[at] = idx_in[i]; }
  }
}

// ---- bounding ball of (up to) 32 points on one warp: Badoiu-Clarkson iterations from the centroid (each step moves
// the centre towards the farthest point by 1/(k+2)); the radius is the exact farthest distance from the final centre
__device__ __forceinline__ float4 warp_ball(float px, float py, float pz, bool valid) {
  const unsigned vm = __ballot_sync(0xffffffffu, valid);
  if (vm == 0u) return make_float4(0.f, 0.f, 0.f, 0.f);
  const bool fin = !valid || (px == px && py == py && pz == pz && fabsf(px) < 1e18f && fabsf(py) < 1e18f && fabsf(pz) < 1e18f);
  if (__ballot_sync(0xffffffffu, !fin) != 0u) return make_float4(0.f, 0.f, 0.f, 1e30f);   // NaN / inf in the group: no culling
  const float inv = 1.0f / (float)__popc(vm);
  float cx = warp_sum(valid ? px : 0.f) * inv, cy = warp_sum(valid ? py : 0.f) * inv, cz = warp_sum(valid ? pz : 0.f) * inv;
  float m = 0.f;
#pragma unroll 1
  for (int it = 0; it <= 10; ++it) {
    const float dx = px - cx, dy = py - cy, dz = pz - cz;
    const float d2 = valid ? fmaf(dx, dx, fmaf(dy, dy, dz * dz)) : -1.f;
    m = __int_as_float(__reduce_max_sync(0xffffffffu, __float_as_int(fmaxf(d2, 0.f))));   // d2 >= 0: the bit pattern is monotone
    if (it == 10) break;
    const int who = __ffs(__ballot_sync(0xffffffffu, valid && d2 >= m)) - 1;
    const float fx = __shfl_sync(0xffffffffu, px, who), fy = __shfl_sync(0xffffffffu, py, who), fz = __shfl_sync(0xffffffffu, pz, who);
    const float s = 1.0f / (float)(it + 2);
    cx = fmaf(fx - cx, s, cx); cy = fmaf(fy - cy, s, cy); cz = fmaf(fz - cz, s, cz);
  }
  return make_float4(cx, cy, cz, sqrtf(m) * (1.f + 2e-6f) + 1e-30f);
}

// one warp per group of the padded layout: sorted points, permutation, group ball
template <class T>
__global__ void __launch_bounds__(256) points_gather_kernel(const T* __restrict__ src, const int* __restrict__ idx_sorted, const FrameTable ft,
                                                            float4* __restrict__ pts, int* __restrict__ perm, float4* __restrict__ grp) {
  const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long span_groups = ft.off[ft.n_frames] >> 5;
  if (g >= span_groups) return;
  int f = 0;
  while (f + 1 < ft.n_frames && (g << 5) >= ft.off[f + 1]) ++f;
  const long long r = (g << 5) - ft.off[f] + lane;        // rank inside the frame
  const bool valid = r < ft.cnt[f];
  float px = 0.f, py = 0.f, pz = 0.f; int id = -1;
  if (valid) {
    id = idx_sorted[ft.src_off[f] + r];
    const float3 p = load_xyz(src, (long long)id);
    px = p.x; py = p.y; pz = p.z;
  }
  const long long o = (g << 5) + lane;
  pts[o] = make_float4(px, py, pz, 0.f);
  if (perm) perm[o] = valid ? (int)(id - ft.src_off[f]) : -1;   // index inside the frame, the caller's order
  const float4 ball = warp_ball(px, py, pz, valid);
  if (lane == 0) grp[g] = ball;
}
// unsorted conversion (clouds too small to be worth sorting, or >= 2^31 points): frame layout, caller's order
template <class T>
__global__ void __launch_bounds__(256) points_convert_kernel(const T* __restrict__ src, const FrameTable ft, float4* __restrict__ pts, float4* __restrict__ grp) {
  const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (g >= (ft.off[ft.n_frames] >> 5)) return;
  int f = 0;
  while (f + 1 < ft.n_frames && (g << 5) >= ft.off[f + 1]) ++f;
  const long long r = (g << 5) - ft.off[f] + lane;
  const bool valid = r < ft.cnt[f];
  float px = 0.f, py = 0.f, pz = 0.f;
  if (valid) { const float3 p = load_xyz(src, ft.src_off[f] + r); px = p.x; py = p.y; pz = p.z; }
  pts[(g << 5) + lane] = make_float4(px, py, pz, 0.f);
  const float4 ball = warp_ball(px, py, pz, valid);
  if (lane == 0) grp[g] = ball;
}
// group balls of a caller-owned float4 buffer (flash_points_set_device: caller's order, one frame)
__global__ void __launch_bounds__(256) groups_ball_kernel(const float4* __restrict__ pts, long long n, float4* __restrict__ grp) {
  const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if ((g << 5) >= n) return;
  const long long i = (g << 5) + lane;
  const bool valid = i < n;
  float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid) p = __ldg(pts + i);
  const float4 ball = warp_ball(p.x, p.y, p.z, valid);
  if (lane == 0) grp[g] = ball;
}

}  // namespace flashk
