// flash_kernels.cuh -- the CUDA kernels of libflashsdf (sm_100a).
//
//   prepare_kernel   one 320-thread block per config: FK, convex world->geometry transforms, RBF control points,
//                    fp64 Gauss-Jordan inverse of the Turk-O'Brien system, one thread per matrix element
//                    (rows A3, A4, A7, A10)
//   cost_kernel      the hot loop (rows A5, A6, A7, A8): points x configs -> per-config accumulators
//                    (cost, per-body wrenches, RBF adjoint sums); also serves skin_eval (DIST mode)
//   epilogue_kernel  adjoint RBF solve + wrench -> joint-space gradient (replaces ForwardDiff's
//                    ceil(nx/10) dual passes with one pass)
//   raycast kernels  batched ray marching of the same field (row A9) + order-preserving compaction
//   iterate_kernel   epilogue + NaiveSolver update + next prepare for the on-device estimate_state loop (row F1)
//   grid kernels     SDF grid baking / error sampling for FLASH_SDF_GRID
//
// Design notes (B200): tensor cores are not used (no dense contraction on this path).  Points are
// read as coalesced float4; the config's transforms and RBF tables are staged in shared memory once
// per block, the hulls' closest-feature BSP trees are read from global memory through L1; reductions
// are warp shuffles -> per-warp shared accumulators -> one fp64 atomicAdd per accumulator per block.
#pragma once
#include <cuda_runtime.h>

#include "flash_dev.h"

namespace flashk {

using namespace flashdev;

#ifndef FLASH_BLOCK
#define FLASH_BLOCK 256
#endif
constexpr int BLOCK = FLASH_BLOCK;
constexpr int WARPS = BLOCK / 32;
// The cost kernel runs ONE 768-thread block per SM (24 warps, 80 registers) whose warps take groups of 32 points
// from a block-wide counter.  History (K5, ms per launch): 256 x 3 static 62.5 -- its ~6000-instruction
// straight-line body was instruction-fetch bound (ncu stall_no_instruction 2.4 per issue) and every block ended
// with 7 warps waiting for the slowest; a block barrier per batch 58.0; 384 x 2 + barrier 56.6 (768 x 1 + barrier
// 62.5: one barrier over 24 warps waits too long); barrier every 2nd batch 55.4; dynamic hand-out of the groups
// with no barrier at all 54.2 (384 x 2), 16128-point tiles 53.5, 768 x 1 52.5 (640 x 1, 96 regs: 53.1; 512 x 1,
// 120 regs, no spills: 55.4; 256 x 3: 57.0).
// Round 2, with the closest-feature BSP narrow phase (no GJK code in the loop: 64 registers, no spills, geometry in L1
// instead of shared memory): 768 x 1 31.2 ms, 384 x 2 30.5, 1024 x 1 28.4, 384 x 3 28.2, 512 x 2 28.1 (K4: 2.94 ms).
#ifndef FLASH_CBLOCK
#define FLASH_CBLOCK 512
#endif
constexpr int CBLOCK = FLASH_CBLOCK;
constexpr int CWARPS = CBLOCK / 32;
#ifndef FLASH_GRAB
#define FLASH_GRAB 2
#endif
constexpr int FRAMES_MAX = 64;   // = MAX_FRAMES of flash_ingest.cuh
constexpr int QENT = 8;  // doubles per skin-queue entry: x y z s ginv nhx nhy nhz (slot in a side array)

struct CostParams {
  const HotModel* hot;
  const F4* geo_f4;
  const uint32_t* geo_u32;
  const float* grid_pool;
  const float4* pts;
  long long n_pts;
  const double* tab;   // [C][tab_len]
  double* acc;         // [C][acc_len]
  int32_t* argmin;     // [C][n_pts] or null
  const int* perm;     // sorted position -> caller index (null: identity)
  const float4* grp;   // per 32 consecutive points: bounding ball {cx, cy, cz, rho} (null: no group culling)
  const uint4* rec;    // group records of group_kernel, [C][rec_stride] (null: per-warp hull cull inside the loop)
  long long rec_stride;
  double* dist;        // [n_pts] (DIST mode, C == 1)
  const long long* frames;   // null: one frame [0, n_pts); else off[0 .. FRAMES_MAX], cnt[FRAMES_MAX + 1 ..] and
  int cfg_per_frame;         // configuration c evaluates frame c / cfg_per_frame only
  int pts_per_block;
  int use_grid;
  double* part;        // deterministic mode: per-tile partial sums [C][gridDim.y][acc_len] (null: fp64 atomics into acc)
  const int* tile_order;   // [frames][gridDim.y] block row -> tile (null: identity): expensive tiles first, so that the kernel does not end on them
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

// dynamic shared memory layout of cost_kernel / raycast_kernel
struct SmemLayout {
  size_t off_hot, off_tab, off_tabf, off_acc, off_queue, off_qslot, total;
};
__host__ __device__ inline SmemLayout smem_layout(int tab_len, int acc_len, int tabf_len, bool grad_skins, int warps = WARPS) {
  SmemLayout L;
  size_t o = 0;
  L.off_hot = o; o = align16(o + sizeof(HotModel));
  L.off_tab = o; o = align16(o + sizeof(double) * (size_t)tab_len);
  L.off_tabf = o; o = align16(o + sizeof(float) * (size_t)(tabf_len > 0 ? tabf_len : 4));
  L.off_acc = o; o = align16(o + sizeof(double) * (size_t)acc_len * warps);
  L.off_queue = o; o = align16(o + (grad_skins ? sizeof(double) * QENT * 32 * warps : 0));
  L.off_qslot = o; o = align16(o + (grad_skins ? sizeof(int) * 32 * warps : 0));
  L.total = o;
  return L;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------------
// One block per config.  Everything that is iterated over (the RBF matrix, the control points) lives in
// shared memory: the serial version of this kernel spent 35 us chasing dependent global loads through
// a 17 x 17 triangular solve.  The Turk-O'Brien system is INVERTED (in-place Gauss-Jordan with partial
// pivoting) and the inverse kept in `aux`: weights and, in the epilogue, the adjoint
// solve are then lane-parallel matrix-vector products (the system is symmetric).
__host__ __device__ inline size_t prepare_smem_doubles(const ModelC& M) {
  size_t need = 8;
  for (int k = 0; k < M.n_skins; ++k) {
    const size_t m = (size_t)M.skin[k].M, n = (size_t)M.skin[k].npts;
    const size_t v = m * m + 3 * n + 3 * m + 4 + (m * m <= 1024 ? m * m : 0);   // A, centres, pivots, pivot row, pivot column, 2 scratch (+2 pad), second matrix of the fast path
    if (v > need) need = v;
  }
  const size_t fk = 25 * MAX_BODIES + 64 + (MAX_BODIES * sizeof(BodyC) + 7) / 8;   // FK staging (reused by the skins afterwards)
  return need > fk ? need : fk;
}
// One BLOCK of PREP_T threads prepares one config: x -> tab, aux; zeroes acc (may be null);
// sm = prepare_smem_doubles() doubles.  The single-warp version spent 18.5 of its 26.5 us in the Gauss-Jordan
// inversion of the 17 x 17 system (clock64 stamps, K3) and 5.5 us in 17 serial fp64 square roots per lane;
// here every matrix element has its own thread: per pivot one warp-shuffle search, the pivot row (scaled) and
// column are copied aside, then all elements update at once -- the same fma per element as before, so the
// inverse is bit-identical.
constexpr int PREP_T = 320;
// 1/x for the pivot candidates of the fast path: MUFU seed (2^-20) + two Newton steps (2^-40, then rounding-limited),
// ~70 cycles instead of the ~350 of the IEEE division routine that every pivot step would wait for
__device__ __forceinline__ double rcp_newton(double x) {
  double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x, y, 1.0); y = fma(y, e, y);
  e = fma(-x, y, 1.0); y = fma(y, e, y);
  return y;
}
__device__ __forceinline__ void prepare_block(const ModelC& M, const double* x, double* tab, double* aux, double* acc, int* status, double* sm, int tid, int T) {
  const int lane = tid & 31, warp = tid >> 5;
#ifdef FLASH_ITER_TIMING
  long long ts[8]; ts[0] = clock64();
#define FLASH_TS(i) ts[i] = clock64()
#else
#define FLASH_TS(i) ((void)0)
#endif
  if (acc) for (int i = tid; i < M.acc_len; i += T) acc[i] = 0.0;   // replaces a memset node
  {
    // forward kinematics is a serial chain: run it on thread 0 out of shared memory (bodies, x and the
    // transforms staged by all threads) instead of chasing dependent global loads
    const int nb = M.n_bodies;
    double* s_aux = sm;                                       // [25 nb]
    double* s_x = s_aux + 25 * MAX_BODIES;                    // [nq <= 64]
    BodyC* s_body = reinterpret_cast<BodyC*>(s_x + 64);       // [nb]
    {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(M.body);
      uint32_t* dst = reinterpret_cast<uint32_t*>(s_body);
      for (int i = tid; i < (int)(nb * sizeof(BodyC) / 4); i += T) dst[i] = src[i];
      for (int i = tid; i < M.nq; i += T) s_x[i] = x[i];
    }
    __syncthreads();
    FLASH_TS(1);
    if (tid == 0) fk_bodies(s_body, nb, M.nq, s_x, s_aux);
    __syncthreads();
    FLASH_TS(2);
    for (int i = tid; i < 25 * nb; i += T) aux[i] = s_aux[i];
    __syncthreads();
  }
  for (int j = tid; j < M.n_convex; j += T) prepare_convex_one(M, j, aux, tab);
  for (int k = 0; k < M.n_skins; ++k) {
    const SkinC& K = M.skin[k];
    const int n = K.npts, Mm = K.M;
    double* A = sm;                 // [Mm][Mm]
    double* cs = A + Mm * Mm;       // [n][3]
    double* piv = cs + 3 * n;       // [Mm]
    double* prow = piv + Mm;        // [Mm] pivot row, scaled
    double* pcol = prow + Mm;       // [Mm] pivot column
    double* scr = pcol + Mm;        // [2] pivot row index, |pivot|
    for (int i = tid; i < n; i += T) {
      const D3 p = skin_centre_one(M, K, i, x, aux);
      cs[3 * i] = p.x; cs[3 * i + 1] = p.y; cs[3 * i + 2] = p.z;
      double* ct = tab + K.tab_off + 3 * i;
      ct[0] = p.x; ct[1] = p.y; ct[2] = p.z;
    }
    __syncthreads();
    const int i0 = tid / Mm, j0 = tid - i0 * Mm, di = T / Mm, dj = T - di * Mm;   // element (i, j) of this thread, and its stride
    for (int e = tid, i = i0, j = j0; e < Mm * Mm; e += T, i += di, j += dj) {       // [Phi Q; Q^T 0]
      if (j >= Mm) { j -= Mm; ++i; }
      double v = 0;
      if (i < n && j < n) {
        const double dx = cs[3 * i] - cs[3 * j], dy = cs[3 * i + 1] - cs[3 * j + 1], dz = cs[3 * i + 2] - cs[3 * j + 2];
        const double r2 = dx * dx + dy * dy + dz * dz;
        v = r2 * sqrt(r2);
      } else if (i < n) v = (j == n) ? 1.0 : cs[3 * i + (j - n - 1)];
      else if (j < n) v = (i == n) ? 1.0 : cs[3 * j + (i - n - 1)];
      A[e] = v;
    }
    __syncthreads();
    FLASH_TS(3);
    bool ok = true;
    if (Mm * Mm <= T && Mm <= 22) {
      // Fast path (K3: 17 x 17 <= 320 threads): every thread owns one element in a register, the matrix in shared
      // memory is only the exchange medium.  Per pivot ONE barrier: the matrix (and the reciprocals) alternate between
      // two buffers, so a step reads the buffer the previous step wrote and writes the other one; the row exchange is done
      // by addressing (row bi "holds" old row kk) instead of a swap phase, and the reciprocal of every pivot candidate was
      // computed by the thread that produced it during the previous update, so no division sits on the critical path.
      // Same update formulas as the generic path below; the reciprocal (rcp_newton, ~1 ulp) and the choice among
      // near-equal pivot candidates (below) differ.
      // (Tried: the whole inversion on ONE warp, a column per lane in registers, multipliers by shuffles, no barrier --
      // 3x slower, 51 k cycles against 17 k: ~400 dependent instructions per pivot on a single warp issue at ~5 cycles each.
      // Tried in track_kernel: Newton-Schulz refinement of the previous iteration's inverse with a residual certificate --
      // never accepted on K3: one solver step moves the system further than |I - A X| < 1/4 allows.)
      // The pivot search is not a scan either: a thread owns element (i, j) with tid = j * Mm + i, so the Mm threads that
      // produce column kk + 1 during step kk sit in one warp (or two), and each such warp finds its largest candidate with
      // ONE redux on a 32-bit key -- the magnitude rounded to fp32 with its low five bits replaced by 31 - row (largest
      // magnitude first, lowest row on ties) -- and leaves it in shared memory; step kk + 1 reads two words.  The pivot is
      // then the largest candidate up to 2^-19 relative, which is all partial pivoting needs (K3: the 17 broadcast loads
      // and dependent fp64 compares per thread and pivot were ~1/3 of the inversion's 16 k cycles).
      const bool act = tid < Mm * Mm;
      const int j = tid / Mm, i = tid - j * Mm;
      const int e = i * Mm + j;                               // where this thread's element lives in the row-major matrices
      const double* cur = A; double* nxt = scr + 4;          // [Mm][Mm] each
      const double* rcp = prow; double* rcn = pcol;          // [Mm] reciprocals of the next pivot column (prow/pcol are unused here)
      unsigned* pk = reinterpret_cast<unsigned*>(scr);        // [2][2] pivot keys of the next column: per buffer, one word per warp that holds it
      unsigned* pkc = pk; unsigned* pkn = pk + 2;
      double a = act ? A[e] : 0.0;
      {
        const bool cand = act && j == 0;
        const unsigned mask = __ballot_sync(0xffffffffu, cand);
        if (cand) {
          const float mg = (float)fabs(a);
          const unsigned key = ((mg == mg ? __float_as_uint(mg) : 0u) & ~31u) | (unsigned)(31 - i);   // NaN: never the pivot
          const unsigned m = __reduce_max_sync(mask, key);
          prow[i] = rcp_newton(a);
          if (lane == __ffs(mask) - 1) { pkc[0] = m; pkc[1] = 0u; }   // column 0 = threads 0 .. Mm-1: one warp
        }
      }
      __syncthreads();
      for (int kk = 0; kk < Mm; ++kk) {
        const unsigned k0 = pkc[0], k1 = pkc[1];
        const unsigned km = k0 > k1 ? k0 : k1;
        const int bi = 31 - (int)(km & 31u);
        if ((km >> 5) == 0u) ok = false;         // every candidate is zero (or below the fp32 range)
        const double inv = rcp[bi];
        if (tid == 0) piv[kk] = (double)bi;
        double nv = 0.0;
        if (act) {
          const double pr = cur[bi * Mm + j];
          const double f = (i == bi) ? cur[kk * Mm + kk] : cur[i * Mm + kk];
          const double own = (i == bi) ? cur[kk * Mm + j] : a;
          if (i == kk) nv = (j == kk) ? inv : pr * inv;
          else nv = (j == kk) ? -f * inv : fma(-f, pr * inv, own);
          nxt[e] = nv; a = nv;
        }
        {   // the candidates of the next pivot: column kk + 1, rows below kk
          const int lo = (kk + 1) * Mm;                                   // first thread of that column
          const bool col = act && j == kk + 1;
          const bool cand = col && i > kk;
          const unsigned mask = __ballot_sync(0xffffffffu, cand);
          if (col) rcn[i] = rcp_newton(nv);
          if (cand) {
            const float mg = (float)fabs(nv);
            const unsigned key = ((mg == mg ? __float_as_uint(mg) : 0u) & ~31u) | (unsigned)(31 - i);
            const unsigned m = __reduce_max_sync(mask, key);
            if (lane == __ffs(mask) - 1) {
              const int slot = warp - ((lo + kk + 1) >> 5);              // warps of the candidates: that of thread lo + kk + 1, and maybe the next
              pkn[slot] = m;
              if (((lo + Mm - 1) >> 5) == ((lo + kk + 1) >> 5)) pkn[1] = 0u;   // all in one warp: the other word is empty
            }
          }
        }
        __syncthreads();                         // the new matrix is complete; nobody reads the old one any more
        { const double* t0 = cur; cur = nxt; nxt = const_cast<double*>(t0); }
        { const double* t1 = rcp; rcp = rcn; rcn = const_cast<double*>(t1); }
        { unsigned* t2 = pkc; pkc = pkn; pkn = t2; }
      }
      FLASH_TS(4);
      // undo the row exchanges on the columns: the exchanges (kk, piv[kk]), kk = Mm-1 .. 0, move this thread's
      // column to `pos`
      int pos = j;
      for (int kk = Mm - 1; kk >= 0; --kk) { const int pv = (int)piv[kk]; pos = (pos == kk) ? pv : ((pos == pv) ? kk : pos); }
      if (act) A[i * Mm + pos] = a;              // a permutation of the columns: no two threads write one address
      __syncthreads();
    } else {
      for (int kk = 0; kk < Mm; ++kk) {
        if (warp == 0) {
          // pivot: largest |A[i][kk]|, i >= kk (lowest row on ties, as the serial factorisation)
          double best = -1.0; int bi = kk;
          for (int i = kk + lane; i < Mm; i += 32) { const double v = fabs(A[i * Mm + kk]); if (v > best) { best = v; bi = i; } }
  #pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
          }
          if (lane == 0) { scr[0] = (double)bi; scr[1] = best; piv[kk] = (double)bi; }
        }
        __syncthreads();
        const int bi = (int)scr[0];
        if (!(scr[1] > 0.0)) ok = false;
        if (bi != kk) {   // block-uniform
          for (int j = tid; j < Mm; j += T) { const double t = A[kk * Mm + j]; A[kk * Mm + j] = A[bi * Mm + j]; A[bi * Mm + j] = t; }
          __syncthreads();
        }
        {
          const double inv = 1.0 / A[kk * Mm + kk];
          for (int j = tid; j < Mm; j += T) { prow[j] = (j == kk) ? inv : A[kk * Mm + j] * inv; pcol[j] = A[j * Mm + kk]; }
          __syncthreads();
          for (int e = tid, i = i0, j = j0; e < Mm * Mm; e += T, i += di, j += dj) {
            if (j >= Mm) { j -= Mm; ++i; }
            const double f = pcol[i];
            A[e] = (i == kk) ? prow[j] : ((j == kk) ? -f * inv : fma(-f, prow[j], A[e]));
          }
          __syncthreads();
        }
      }
      FLASH_TS(4);
      for (int kk = Mm - 1; kk >= 0; --kk) {      // undo the row exchanges on the columns
        const int pv = (int)piv[kk];
        if (pv != kk) {   // block-uniform
          for (int i = tid; i < Mm; i += T) { const double t = A[i * Mm + kk]; A[i * Mm + kk] = A[i * Mm + pv]; A[i * Mm + pv] = t; }
          __syncthreads();
        }
      }
    }
    FLASH_TS(5);
    if (!ok && tid == 0) atomicExch(status, 1);   // singular RBF system (coincident control points)
    // weights: rhs = 0 (surface), -1 (skeleton) (src/Flash.jl:211)  ->  w = -sum_{skeleton j} Ainv[:, j]
    double* w = tab + K.tab_off + 3 * n;
    for (int i = tid; i < Mm; i += T) {
      double sacc = 0;
      for (int j = K.nsurf; j < n; ++j) sacc -= A[i * Mm + j];
      w[i] = sacc;
    }
    if (Mm != n + 4) for (int i = tid; i < 3; i += T) w[n + 1 + i] = 0.0;
    double* Ainv = aux + aux_skin0(M) + K.aux_off;
    for (int e = tid; e < Mm * Mm; e += T) Ainv[e] = A[e];
    __syncthreads();
    FLASH_TS(6);
#ifdef FLASH_ITER_TIMING
    if (tid == 0 && blockIdx.x == 0) printf("prepare_block cycles: stage %lld fk %lld centres+assemble %lld invert %lld unswap %lld store %lld\n", ts[1] - ts[0], ts[2] - ts[1], ts[3] - ts[2], ts[4] - ts[3], ts[5] - ts[4], ts[6] - ts[5]);
#endif
  }
}

__global__ void __launch_bounds__(PREP_T) prepare_kernel(const ModelC* __restrict__ Mp, const double* __restrict__ X, double* __restrict__ tab_all,
                                                      double* __restrict__ aux_all, double* __restrict__ acc_all, int* __restrict__ status) {
  extern __shared__ __align__(16) unsigned char smem[];
  const ModelC& M = *Mp;
  const int c = blockIdx.x;
  prepare_block(M, X + (size_t)c * M.nx, tab_all + (size_t)c * M.tab_len, aux_all + (size_t)c * M.aux_len,
                acc_all ? acc_all + (size_t)c * M.acc_len : nullptr, status, reinterpret_cast<double*>(smem), threadIdx.x, PREP_T);
}

// ------------------------------------------------------------------------------------------------
// stage the model and this config's tables in shared memory
__device__ __forceinline__ void stage_block(unsigned char* smem, const SmemLayout& L, const HotModel* hot, const double* tab, int tab_len, int nthreads = BLOCK) {
  const int tid = threadIdx.x;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(hot);
    uint32_t* dst = reinterpret_cast<uint32_t*>(smem + L.off_hot);
    for (int i = tid; i < (int)(sizeof(HotModel) / 4); i += nthreads) dst[i] = src[i];
  }
  {
    double* dst = reinterpret_cast<double*>(smem + L.off_tab);
    for (int i = tid; i < tab_len; i += nthreads) dst[i] = tab[i];
    // fp32 side table: hull transforms (culling bounds) and the skins' pre-filter tables
    tabf_fill(*hot, tab, reinterpret_cast<float*>(smem + L.off_tabf), tid, nthreads);
  }
}

// ------------------------------------------------------------------------------------------------
// Group pre-pass as its own kernel: ONE THREAD per (configuration, group of 32 points).  Everything it decides is
// decided once for 32 evaluations (flash_dev.h group_prepass): the candidate hulls from exact distances at the ball
// centre, the nearest one, the BSP nodes from which the group's points finish their descent, the skins that cannot be
// the minimum and an upper bound of the group's minimum.  Inside cost_kernel (round 1: a phase in front of the point
// loop, one 768-thread block per SM) the same code left two thirds of the block waiting at a barrier behind long
// serial threads; here it runs at full occupancy and the point loop only loads a 16-byte record per group:
//   x = cand (bits 0-23) | skin skip bits << 24      y = cut      z, w = starts (3 x 16 bits) | first << 16 in w
struct GroupParams {
  const HotModel* hot;
  const F4* geo_f4;
  const uint32_t* geo_u32;
  const double* tab;       // [C][tab_len]
  const float4* grp;       // group balls
  const long long* frames; // as CostParams
  int cfg_per_frame;
  long long n_groups;      // groups of the (single) frame when frames == null
  uint4* rec;              // [C][rec_stride]
  long long rec_stride;
  int groups_per_block;
  int exact, want_cull;
  unsigned* tile_w;        // [frames][tiles] work estimate of every tile of cost_kernel, summed over the frame's configurations (null: not wanted)
  int groups_per_tile, tiles;
};
constexpr int GBLOCK = 256;
__host__ __device__ inline size_t group_smem(int tabf_len) { return align16(sizeof(HotModel)) + align16(sizeof(float) * (size_t)(tabf_len > 0 ? tabf_len : 4)); }
__global__ void __launch_bounds__(GBLOCK, 4) group_kernel(const GroupParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int tid = threadIdx.x, c = blockIdx.x;
  long long g0 = 0, g1 = p.n_groups;
  int frame = 0;
  if (p.frames) { frame = c / p.cfg_per_frame; g0 = p.frames[frame] >> 5; g1 = g0 + ((p.frames[FRAMES_MAX + 1 + frame] + 31) >> 5); }
  const long long b0 = g0 + (long long)blockIdx.y * p.groups_per_block;
  if (b0 >= g1) return;
  long long b1 = b0 + p.groups_per_block; if (b1 > g1) b1 = g1;
  const HotModel* hot_g = p.hot;
  HotModel* Hs = reinterpret_cast<HotModel*>(smem);
  float* tabf = reinterpret_cast<float*>(smem + align16(sizeof(HotModel)));
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(hot_g);
    uint32_t* dst = reinterpret_cast<uint32_t*>(Hs);
    for (int i = tid; i < (int)(sizeof(HotModel) / 4); i += GBLOCK) dst[i] = src[i];
    tabf_fill(*hot_g, p.tab + (size_t)c * hot_g->tab_len, tabf, tid, GBLOCK);
  }
  __syncthreads();
  const HotModel& H = *Hs;
  GeoView G; G.f4 = p.geo_f4; G.u32 = p.geo_u32; G.grid_pool = nullptr;
  uint4* out = p.rec + (size_t)c * p.rec_stride - g0;
  const unsigned all_skins = H.n_skins >= 32 ? 0xffffffffu : ((1u << H.n_skins) - 1u);
  for (long long base = b0; base < b1; base += GBLOCK) {   // block-uniform trip count: the warps vote below
    const long long g = base + tid;
    const bool on = g < b1;
    unsigned w = 0u;
    if (on) {
      const float4 gb = __ldg(p.grp + g);
      const GroupPre gp = group_prepass(H, G, tabf, gb.x, gb.y, gb.z, gb.w, p.exact != 0, p.want_cull != 0);
      out[g] = make_uint4((gp.cand & 0xffffffu) | (gp.skip << 24), __float_as_uint(gp.cut), (unsigned)(gp.starts & 0xffffffffull),
                          (unsigned)((gp.starts >> 32) & 0xffffull) | ((unsigned)(gp.first < 0 ? 255 : gp.first) << 16));
      // work estimate of the group in cost_kernel: one descent per candidate hull, several times that per skin still in play
      w = 1u + (unsigned)__popc(gp.cand & 0xffffffu) + 6u * (unsigned)__popc(all_skins & ~gp.skip);
    }
    if (p.tile_w) {   // (block-uniform) consecutive lanes hold consecutive groups: one atomic per tile and warp
      const int tl = on ? (int)((g - g0) / p.groups_per_tile) : -1;
      const unsigned peers = __match_any_sync(0xffffffffu, tl);
      const unsigned sum = __reduce_add_sync(peers, w);
      if (on && (__ffs(peers) - 1) == (tid & 31)) atomicAdd(p.tile_w + (size_t)frame * p.tiles + tl, sum);
    }
  }
}

// Order of the tiles of cost_kernel: heaviest first (longest-processing-time-first), from the estimates of group_kernel.
// The cloud is in Hilbert order and its last tiles are often the expensive ones (near the bodies); launched last, their
// blocks run on while most SMs have gone idle (K5 shard of 1.25 M points: 2.62 ms in Hilbert order, 2.50 ms reversed).
// One block per frame of blockDim.x = the power of two >= tiles (<= 1024) threads.  Leaves the estimates zeroed for the next launch.
__global__ void __launch_bounds__(1024) tile_order_kernel(unsigned* __restrict__ tile_w, int tiles, int* __restrict__ order) {
  __shared__ unsigned long long key[1024];
  const int t = threadIdx.x;
  tile_w += (size_t)blockIdx.x * tiles; order += (size_t)blockIdx.x * tiles;   // one block per frame
  key[t] = t < tiles ? (((unsigned long long)tile_w[t] << 32) | (unsigned)(0xffffffffu - (unsigned)t)) : 0ull;
  if (t < tiles) tile_w[t] = 0u;
  __syncthreads();
  for (int k = 2; k <= (int)blockDim.x; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      const int x = t ^ j;
      if (x > t) {
        const unsigned long long a = key[t], b = key[x];
        const bool desc = (t & k) == 0;       // descending overall
        if (desc ? a < b : a > b) { key[t] = b; key[x] = a; }
      }
      __syncthreads();
    }
  if (t < tiles) order[t] = (int)(0xffffffffu - (unsigned)(key[t] & 0xffffffffull));
}

// GRAD: accumulate gradient terms; ARGMIN: write the argmin surface per point; DIST: write the
// min distance per point (skin_eval entry) and skip all accumulation.
#ifndef FLASH_MINB
#define FLASH_MINB 2
#endif
// The point loop of cost_kernel (and of track_kernel): the warps of one block work through the groups of the tile
// [tile0, tile1) of configuration c and add into their own accumulators acc_w; returns this lane's part of the cost.
template <bool GRAD, bool ARGMIN, bool DIST, bool DET, int GRAB = FLASH_GRAB, bool ILP2 = false>
__device__ __forceinline__ double cost_groups(const CostParams& p, const HotModel& H, const GeoView& G, const double* tab, const float* tabf, double* acc_w,
                                              double* queue, int* qslot, int* s_next, int c, int n_convex, long long range0, long long tile0, long long tile1,
                                              int lane, int warp) {
  const unsigned all_convex = n_convex >= 32 ? 0xffffffffu : ((1u << n_convex) - 1u);
  double cost_local = 0.0;

  // Groups of 32 points are handed to the block's warps dynamically (FLASH_GRAB consecutive groups per shared-memory
  // atomic): the cost per group varies several-fold (candidate hulls, skin winners) and with a static schedule every
  // block ended with most warps waiting for the slowest one.  Deterministic mode
  // (flash_options.deterministic) uses the static schedule -- warp w takes chunks w, w + CWARPS, ... -- so that every
  // accumulator sees its terms in one fixed order.
  const int ngroups = (int)((tile1 - tile0 + 31) >> 5);
  // group records (group_kernel, run before this kernel; layout there)
  const bool grp_pre = !DIST && p.rec != nullptr;
  const uint4* rec = grp_pre ? p.rec + (size_t)c * p.rec_stride + ((tile0 - range0) >> 5) : nullptr;
  // chunks of FLASH_GRAB (a power of two) groups start at multiples of FLASH_GRAB: the end of a chunk is a bit test
  int g = 0;
  if (DET) g = warp * GRAB;
  else {
    if (lane == 0) g = atomicAdd(s_next, GRAB);
    g = __shfl_sync(0xffffffffu, g, 0);
  }
  while (g < ngroups) {
    const long long wbase = tile0 + (long long)g * 32;
    const int gcur = g;
    if (((++g) & (GRAB - 1)) == 0) {
      if (DET) g += (CWARPS - 1) * GRAB;
      else {
        if (lane == 0) g = atomicAdd(s_next, GRAB);
        g = __shfl_sync(0xffffffffu, g, 0);
      }
    }
    const long long i = wbase + lane;
    const bool valid = i < tile1;
    PointEval e;
    e.d = 0.f; e.arg = -1; e.slot = -1; e.is_skin = 0;
    e.y = f3(0.f, 0.f, 0.f); e.n = f3(0.f, 0.f, 0.f);   // lanes past the tile end must contribute exact zeros
    float px = 0.f, py = 0.f, pz = 0.f;
    // hulls that can be the minimum for some point of this warp's 32 neighbours, from the group's record
    unsigned cand = all_convex; int first = -1;
    unsigned skin_skip = 0u; float cut0 = 3.0e38f; unsigned long long starts = 0ull;
    if (grp_pre) {
      const uint4 r = __ldg(rec + gcur);
      cand = r.x & 0xffffffu; skin_skip = r.x >> 24; cut0 = __uint_as_float(r.y);
      starts = (unsigned long long)r.z | ((unsigned long long)(r.w & 0xffffu) << 32);
      first = (r.w >> 16) == 255u ? -1 : (int)(r.w >> 16);
    }
    if (valid) {
      const float4 P = __ldg(p.pts + i);
      px = P.x; py = P.y; pz = P.z;
      eval_point(H, G, tab, tabf, (double)px, (double)py, (double)pz, cand, first, p.use_grid != 0, true, e, skin_skip, cut0, starts);
    }
    if (DIST) {
      if (valid) {
        p.dist[i] = e.is_skin ? e.sk.s : (double)e.d;
        if (ARGMIN) p.argmin[i] = e.arg;
      }
      continue;
    }
    if (ARGMIN && valid) p.argmin[(size_t)c * p.n_pts + (p.perm ? (long long)p.perm[i] : i)] = e.arg;
    if (valid) { const double dv = e.is_skin ? e.sk.s : (double)e.d; cost_local = fma(dv, dv, cost_local); }
    if (!GRAD) continue;

    // ---- convex winners: per-body wrench in the geometry frame, g = -2 d n, tau = y x g
    const bool isc = valid && !e.is_skin && e.slot >= 0;
    unsigned remaining = __ballot_sync(0xffffffffu, isc);
    while (remaining) {
      const int leader = __ffs(remaining) - 1;
      const int j = __shfl_sync(0xffffffffu, e.slot, leader);
      const bool mine = isc && e.slot == j;
      const unsigned m = __ballot_sync(0xffffffffu, mine);
      const float g = -2.f * e.d;
      float v0 = mine ? g * e.n.x : 0.f, v1 = mine ? g * e.n.y : 0.f, v2 = mine ? g * e.n.z : 0.f;
      float t0 = mine ? e.y.y * v2 - e.y.z * v1 : 0.f, t1 = mine ? e.y.z * v0 - e.y.x * v2 : 0.f, t2 = mine ? e.y.x * v1 - e.y.y * v0 : 0.f;
#ifndef FLASH_NO_BUTTERFLY
      {
        // all six sums in ONE butterfly: each exchange step halves the number of values a lane still carries
        // (4 + 2 + 1 + 1 + 1 = 9 shuffles instead of 6 x 5); lanes 4k .. 4k+3 end up with the total of value k
        const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
        float k0 = b4 ? t1 : v0, k1 = b4 ? t2 : v1, k2 = b4 ? 0.f : v2, k3 = b4 ? 0.f : t0;      // keep {v0,v1,v2,t0} | {t1,t2,0,0}
        const float s0 = b4 ? v0 : t1, s1 = b4 ? v1 : t2, s2 = b4 ? v2 : 0.f, s3 = b4 ? t0 : 0.f;
        k0 += __shfl_xor_sync(0xffffffffu, s0, 16); k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
        k2 += __shfl_xor_sync(0xffffffffu, s2, 16); k3 += __shfl_xor_sync(0xffffffffu, s3, 16);
        float h0 = b3 ? k2 : k0, h1 = b3 ? k3 : k1;
        h0 += __shfl_xor_sync(0xffffffffu, b3 ? k0 : k2, 8); h1 += __shfl_xor_sync(0xffffffffu, b3 ? k1 : k3, 8);
        float w = b2 ? h1 : h0;
        w += __shfl_xor_sync(0xffffffffu, b2 ? h0 : h1, 4);
        w += __shfl_xor_sync(0xffffffffu, w, 2);
        w += __shfl_xor_sync(0xffffffffu, w, 1);
        // value index of this lane: b4 ? (4,5,-,-) : (0,1,2,3), entry 2*b3 + b2
        const int vi = (b4 ? 4 : 0) + (b3 ? 2 : 0) + (b2 ? 1 : 0);
        if ((lane & 3) == 0 && vi < 6) acc_w[1 + 6 * j + vi] += (double)w;
      }
#else
      v0 = warp_sum(v0); v1 = warp_sum(v1); v2 = warp_sum(v2);
      t0 = warp_sum(t0); t1 = warp_sum(t1); t2 = warp_sum(t2);
      if (lane == 0) {
        double* a = acc_w + 1 + 6 * j;
        a[0] += (double)v0; a[1] += (double)v1; a[2] += (double)v2;
        a[3] += (double)t0; a[4] += (double)t1; a[5] += (double)t2;
      }
#endif
      remaining &= ~m;
    }

    // ---- skin winners: queue them per warp, then lanes = control points (no shuffles)
    if (H.n_skins > 0) {
      const bool iss = valid && e.is_skin;
      const unsigned ms = __ballot_sync(0xffffffffu, iss);
      if (ms) {
        if (iss) {
          const int pos = __popc(ms & ((1u << lane) - 1u));
          double* q = queue + pos * QENT;
          q[0] = (double)px; q[1] = (double)py; q[2] = (double)pz;
          q[3] = e.sk.s; q[4] = e.sk.ginv; q[5] = e.sk.nh.x; q[6] = e.sk.nh.y; q[7] = e.sk.nh.z;
          qslot[pos] = e.slot;
        }
        __syncwarp();
        const int cnt = __popc(ms);
        for (int k = 0; k < H.n_skins; ++k) {
          const SkinC& K = H.skin[k];
          const double* cc = tab + K.tab_off;
          const int n = K.npts;
          // sub-warps of G lanes (G = smallest power of two >= n+1, at most 32) each take one queued
          // point at a time; lane % G is the control point (n = the polynomial pseudo-centre)
          int G = 32;
          while ((G >> 1) > n) G >>= 1;
          const int ci0 = lane & (G - 1), sub = lane / G, E = 32 / G;
          for (int cb = 0; cb <= n; cb += G) {
            const int ci = cb + ci0;
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            if (ci <= n) {
              if (ILP2) {
                // two queue entries per trip in straight-line code (track_kernel: a handful of warps per SM, nothing else
                // to hide the fp64 latency behind); same terms added in the same order
                for (int qe = sub; qe < cnt; qe += 2 * E) {
                  const int qb = qe + E < cnt ? qe + E : qe;
                  const bool u0 = qslot[qe] == k, u1 = qe + E < cnt && qslot[qb] == k;
                  const double* q = queue + qe * QENT;
                  const double* r = queue + qb * QENT;
                  double o[4], o2[4];
                  skin_adjoint_term(cc, cc + 3 * n, n, H.normalize != 0, ci, q[0], q[1], q[2], q[3], q[4], d3(q[5], q[6], q[7]), o);
                  skin_adjoint_term(cc, cc + 3 * n, n, H.normalize != 0, ci, r[0], r[1], r[2], r[3], r[4], d3(r[5], r[6], r[7]), o2);
                  if (u0) { a0 += o[0]; a1 += o[1]; a2 += o[2]; a3 += o[3]; }
                  if (u1) { a0 += o2[0]; a1 += o2[1]; a2 += o2[2]; a3 += o2[3]; }
                }
              } else {
                for (int qe = sub; qe < cnt; qe += E) {
                  if (qslot[qe] != k) continue;
                  const double* q = queue + qe * QENT;
                  double o[4];
                  skin_adjoint_term(cc, cc + 3 * n, n, H.normalize != 0, ci, q[0], q[1], q[2], q[3], q[4], d3(q[5], q[6], q[7]), o);
                  a0 += o[0]; a1 += o[1]; a2 += o[2]; a3 += o[3];
                }
              }
            }
            for (int o = G; o < 32; o <<= 1) {
              a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o);
              a2 += __shfl_xor_sync(0xffffffffu, a2, o); a3 += __shfl_xor_sync(0xffffffffu, a3, o);
            }
            if (sub == 0 && ci <= n) {
              double* gY = acc_w + K.acc_off;
              if (ci < n) { gY[ci] += a0; double* gC = gY + n + 4 + 3 * ci; gC[0] += a1; gC[1] += a2; gC[2] += a3; }
              else { gY[n] += a0; gY[n + 1] += a1; gY[n + 2] += a2; gY[n + 3] += a3; }
            }
          }
        }
        __syncwarp();
      }
    }
  }
  return cost_local;
}

constexpr size_t COST_STATIC_SMEM = 64;   // static __shared__ of cost_kernel (the group hand-out counter)
// DET: deterministic mode as a template parameter -- as a run-time branch on p.part it cost the default kernel 6 % (the
// initial `g = warp` keeps one more value live across the prologue: 104 B of spills at the 64-register cap).
template <bool GRAD, bool ARGMIN, bool DIST, bool DET = false>
__global__ void __launch_bounds__(CBLOCK, FLASH_MINB) cost_kernel(const CostParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int c = blockIdx.x;
  long long range0 = 0, range1 = p.n_pts;
  int frame = 0;
  if (p.frames) { frame = c / p.cfg_per_frame; range0 = p.frames[frame]; range1 = range0 + p.frames[FRAMES_MAX + 1 + frame]; }
  const int ty = p.tile_order ? __ldg(p.tile_order + (size_t)frame * gridDim.y + blockIdx.y) : (int)blockIdx.y;   // one order per frame
  if (range0 + (long long)ty * p.pts_per_block >= range1) return;   // (block-uniform) a frame shorter than the longest one
  // every block re-derives the layout from the same numbers the host used
  const HotModel* hot_g = p.hot;
  const int tab_len = hot_g->tab_len, acc_len = hot_g->acc_len, n_convex = hot_g->n_convex;
  const SmemLayout L = smem_layout(tab_len, acc_len, hot_g->tabf_len, GRAD && hot_g->n_skins > 0, CWARPS);
  stage_block(smem, L, hot_g, p.tab + (size_t)c * tab_len, tab_len, CBLOCK);
  double* acc_w = reinterpret_cast<double*>(smem + L.off_acc) + (size_t)warp * acc_len;
  __shared__ int s_next;
  if (tid == 0) s_next = 0;
  if (!DIST) for (int i = lane; i < acc_len; i += 32) acc_w[i] = 0.0;
  __syncthreads();
  const HotModel& H = *reinterpret_cast<const HotModel*>(smem + L.off_hot);
  GeoView G;
  G.f4 = p.geo_f4;
  G.u32 = p.geo_u32;
  G.grid_pool = p.grid_pool;
  const double* tab = reinterpret_cast<const double*>(smem + L.off_tab);
  const float* tabf = reinterpret_cast<const float*>(smem + L.off_tabf);
  double* queue = reinterpret_cast<double*>(smem + L.off_queue) + (size_t)warp * 32 * QENT;
  int* qslot = reinterpret_cast<int*>(smem + L.off_qslot) + warp * 32;

  const long long tile0 = range0 + (long long)ty * p.pts_per_block;
  long long tile1 = tile0 + p.pts_per_block;
  if (tile1 > range1) tile1 = range1;
  double cost_local = cost_groups<GRAD, ARGMIN, DIST, DET>(p, H, G, tab, tabf, acc_w, queue, qslot, &s_next, c, n_convex, range0, tile0, tile1, lane, warp);
  if (DIST) return;
  cost_local = warp_sum(cost_local);
  if (lane == 0) acc_w[0] += cost_local;
  __syncthreads();
  const double* acc_all = reinterpret_cast<const double*>(smem + L.off_acc);
  const int nacc = GRAD ? acc_len : 1;
  for (int l = tid; l < nacc; l += CBLOCK) {
    double s = 0;
#pragma unroll
    for (int w = 0; w < CWARPS; ++w) s += acc_all[(size_t)w * acc_len + l];
    if (DET) p.part[((size_t)c * gridDim.y + ty) * acc_len + l] = s;
    else if (s != 0.0) atomicAdd(p.acc + (size_t)c * acc_len + l, s);
  }
}

// deterministic mode: acc[c][l] = sum over the tiles of configuration c, in tile order
__global__ void __launch_bounds__(128) reduce_partials_kernel(const double* __restrict__ part, double* __restrict__ acc, int acc_len, int tiles_max,
                                                              const long long* __restrict__ frames, int cfg_per_frame, long long n_pts, int pts_per_block, int nacc) {
  const int c = blockIdx.x;
  long long cnt = n_pts;
  if (frames) cnt = frames[FRAMES_MAX + 1 + c / cfg_per_frame];
  const int tiles = (int)((cnt + pts_per_block - 1) / pts_per_block);
  for (int l = threadIdx.x; l < nacc; l += 128) {
    double s = 0.0;
    for (int t = 0; t < tiles && t < tiles_max; ++t) s += part[((size_t)c * tiles_max + t) * acc_len + l];
    acc[(size_t)c * acc_len + l] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// One warp per config (after the all-reduce): the warp-parallel form of flashdev::epilogue_config, same
// formulas.  lambda = A^-1 gY is a matrix-vector product with the inverse prepare_kernel kept (the
// system is symmetric); control points, hulls and joints are spread over the lanes; per-body wrenches
// are accumulated in shared memory.
__host__ __device__ inline size_t epilogue_smem_doubles(const ModelC& M) {
  size_t need = 8;
  for (int k = 0; k < M.n_skins; ++k) {
    const size_t m = (size_t)M.skin[k].M, n = (size_t)M.skin[k].npts;
    const size_t v = m * m + 3 * n + 2 * m + 4;
    if (v > need) need = v;
  }
  return need + 6 * MAX_BODIES;
}
// one warp finishes one config: acc (summed over points and ranks) -> *cost_c and grad (may be null);
// sm = epilogue_smem_doubles() doubles
// preloaded: the caller's whole block has already copied skin 0's inverse and table into `sm` (epilogue_preload)
__device__ __forceinline__ void epilogue_preload(const ModelC& M, const double* tab, const double* aux, double* sm, int tid, int T) {
  if (M.n_skins < 1) return;
  const SkinC& K = M.skin[0];
  const int n = K.npts, Mm = K.M;
  double* Ainv = sm + 6 * MAX_BODIES;
  double* cc = Ainv + Mm * Mm;
  const double* src = aux + aux_skin0(M) + K.aux_off;
  for (int e = tid; e < Mm * Mm; e += T) Ainv[e] = src[e];
  const double* ct = tab + K.tab_off;
  for (int e = tid; e < 4 * n + 4; e += T) cc[e] = ct[e];
}
__device__ __forceinline__ void epilogue_warp(const ModelC& M, const double* x, const double* tab, const double* aux, const double* acc, double w_def,
                                              double* cost_c, double* grad, double* sm, int lane, bool preloaded = false, bool direct = false) {
  double* F = sm;                                // [nb][3]
  double* Mo = F + 3 * MAX_BODIES;               // [nb][3]
  double* sk = Mo + 3 * MAX_BODIES;              // per-skin scratch
  const int nb = M.n_bodies;
#ifdef FLASH_ITER_TIMING
  long long es[8]; for (int q = 0; q < 8; ++q) es[q] = 0; es[0] = clock64();
#define FLASH_ES(i) es[i] = clock64()
#else
#define FLASH_ES(i) ((void)0)
#endif
  if (grad) {
    for (int i = lane; i < 6 * MAX_BODIES; i += 32) F[i] = 0.0;
    for (int k = lane; k < M.nx; k += 32) grad[k] = 0.0;
    __syncwarp();
    // convex wrenches, accumulated in the geometry frame: f = sum g, tau = sum y x g, g = -2 d n
    for (int j = lane; j < M.n_convex; j += 32) {
      const ConvexC& Cv = M.cvx[j];
      if (Cv.body < 0) continue;
      double S[12];
      xf_compose(aux + aux_X(M, Cv.body), Cv.frame, S);
      const double* a = acc + 1 + 6 * j;
      const D3 f = xf_rot(S, d3(a[0], a[1], a[2]));
      const D3 m = xf_rot(S, d3(a[3], a[4], a[5])) + cross(xf_t(S), f);
      atomicAdd(F + 3 * Cv.body, f.x); atomicAdd(F + 3 * Cv.body + 1, f.y); atomicAdd(F + 3 * Cv.body + 2, f.z);
      atomicAdd(Mo + 3 * Cv.body, m.x); atomicAdd(Mo + 3 * Cv.body + 1, m.y); atomicAdd(Mo + 3 * Cv.body + 2, m.z);
    }
    for (int k = 0; k < M.n_skins; ++k) {
      const SkinC& K = M.skin[k];
      const int n = K.npts, Mm = K.M;
      // direct: tab and aux are already in shared memory (track_kernel) -- no copy, only lam needs scratch
      const double* Ainv = direct ? aux + aux_skin0(M) + K.aux_off : sk;   // [Mm][Mm]
      const double* cc = direct ? tab + K.tab_off : sk + Mm * Mm;          // [n][3]
      const double* w = cc + 3 * n;                                         // [n + 4]
      double* lam = direct ? sk : sk + Mm * Mm + 4 * n + 4;                 // [Mm]
      const double* gY = acc + K.acc_off;
      const double* gC = gY + n + 4;
      __syncwarp();
      if (!direct && !(preloaded && k == 0)) {
        const double* src = aux + aux_skin0(M) + K.aux_off;
        for (int e = lane; e < Mm * Mm; e += 32) sk[e] = src[e];
        const double* ct = tab + K.tab_off;
        for (int e = lane; e < 4 * n + 4; e += 32) sk[Mm * Mm + e] = ct[e];   // centres then (w, a0, a): contiguous in tab
      }
      __syncwarp();
      FLASH_ES(1);
      for (int i = lane; i < Mm; i += 32) {
        double sacc = 0;
        for (int j = 0; j < Mm; ++j) sacc = fma(Ainv[i * Mm + j], gY[j], sacc);
        lam[i] = sacc;
      }
      __syncwarp();
      FLASH_ES(2);
      const D3 la = Mm == n + 4 ? d3(lam[n + 1], lam[n + 2], lam[n + 3]) : d3(0, 0, 0);
      const D3 av = d3(w[n + 1], w[n + 2], w[n + 3]);
      for (int i0 = 0; i0 < n; i0 += 32) {          // warp-uniform trip count: the wrench sums below are warp reductions
        const int i = i0 + lane;
        const bool on = i < n;
        D3 ci = d3(0, 0, 0), Gi = d3(0, 0, 0);
        int b = -1;
        if (on) {
          ci = d3(cc[3 * i], cc[3 * i + 1], cc[3 * i + 2]);
          Gi = d3(gC[3 * i], gC[3 * i + 1], gC[3 * i + 2]);
          for (int j = 0; j < n; ++j) {
            const D3 u = ci - d3(cc[3 * j], cc[3 * j + 1], cc[3 * j + 2]);
            const double r = dsqrt_fast(dot(u, u));   // MUFU seed + Newton (8.5e-14 relative; 0 -> 0): the IEEE sqrt cost 400 cycles x n per lane
            Gi = Gi - u * (3.0 * r * (lam[i] * w[j] + lam[j] * w[i]));
          }
          Gi = Gi - av * lam[i] - la * w[i];
          b = M.skin_pt_body[K.pt_off + i];
          if (K.kind == SURF_DEFORMABLE_SKIN && i < K.nsurf) {
            const D3 gd = b >= 0 ? xf_rotT(aux + aux_X(M, b), Gi) : Gi;
            double* o = grad + M.nq + K.deform_off + 3 * i;
            o[0] += gd.x; o[1] += gd.y; o[2] += gd.z;
          }
        }
        FLASH_ES(3);
        // per distinct body among the lanes: one shuffle reduction of (force, moment) instead of fp64 shared
        // atomics (CAS loops; 13 lanes on one address cost 3 us of the 12.5 us epilogue on K3)
        unsigned rest = __ballot_sync(0xffffffffu, on && b >= 0);
        while (rest) {
          const int bb = __shfl_sync(0xffffffffu, b, __ffs(rest) - 1);
          const bool mine = on && b == bb;
          const D3 m = cross(ci, Gi);
          const double fx = warp_sum(mine ? Gi.x : 0.0), fy = warp_sum(mine ? Gi.y : 0.0), fz = warp_sum(mine ? Gi.z : 0.0);
          const double mx = warp_sum(mine ? m.x : 0.0), my = warp_sum(mine ? m.y : 0.0), mz = warp_sum(mine ? m.z : 0.0);
          if (lane == 0) { F[3 * bb] += fx; F[3 * bb + 1] += fy; F[3 * bb + 2] += fz; Mo[3 * bb] += mx; Mo[3 * bb + 1] += my; Mo[3 * bb + 2] += mz; }
          rest &= ~__ballot_sync(0xffffffffu, mine);
        }
        __syncwarp();
      }
    }
    __syncwarp();
    FLASH_ES(4);
    if (lane == 0)
      for (int b = nb - 1; b >= 0; --b) {  // subtree wrenches: children follow parents
        const int p = M.body[b].parent;
        if (p >= 0) for (int k = 0; k < 3; ++k) { F[3 * p + k] += F[3 * b + k]; Mo[3 * p + k] += Mo[3 * b + k]; }
      }
    __syncwarp();
    for (int b = lane; b < nb; b += 32) {
      const BodyC& B = M.body[b];
      if (B.jtype == JOINT_FIXED) continue;
      const double* P = aux + aux_P(M, b);
      const D3 Fb = d3(F[3 * b], F[3 * b + 1], F[3 * b + 2]), Mb = d3(Mo[3 * b], Mo[3 * b + 1], Mo[3 * b + 2]);
      if (B.jtype == JOINT_REVOLUTE) {
        const D3 wv = xf_rot(P, d3(B.axis[0], B.axis[1], B.axis[2]));
        grad[B.qoff] += dot(wv, Mb - cross(xf_t(P), Fb));
      } else {
        const double* Xb = aux + aux_X(M, b);
        const D3 Gp = xf_rotT(P, Mb - cross(xf_t(Xb), Fb));
        const D3 Fp = xf_rotT(P, Fb);
        const double nrm = aux[aux_qnorm(M, b)];
        const double* q = x + B.qoff;
        const double rn = 1.0 / nrm;
        const double qw = q[0] * rn;
        const D3 v = d3(q[1] * rn, q[2] * rn, q[3] * rn);
        const D3 gv = Gp * qw + cross(Gp, v);
        const double inv = 2.0 * rn;
        grad[B.qoff] += -dot(Gp, v) * inv;
        grad[B.qoff + 1] += gv.x * inv; grad[B.qoff + 2] += gv.y * inv; grad[B.qoff + 3] += gv.z * inv;
        grad[B.qoff + 4] += Fp.x; grad[B.qoff + 5] += Fp.y; grad[B.qoff + 6] += Fp.z;
      }
    }
    __syncwarp();
  }
  FLASH_ES(5);
  double reg = 0;
  for (int k = M.nq + lane; k < M.nx; k += 32) { reg = fma(x[k], x[k], reg); if (grad) grad[k] += 2.0 * w_def * x[k]; }
  reg = warp_sum(reg);
  if (lane == 0) *cost_c = acc[0] + w_def * reg;
#ifdef FLASH_ITER_TIMING
  if (lane == 0 && blockIdx.x == 0 && grad) printf("epilogue cycles: load %lld lambda %lld Gloop %lld reduce %lld bodies %lld reg %lld\n", es[1] - es[0], es[2] - es[1], es[3] - es[2], es[4] - es[3], es[5] - es[4], clock64() - es[5]);
#endif
}
__global__ void __launch_bounds__(32) epilogue_kernel(const ModelC* __restrict__ Mp, const double* __restrict__ X, const double* __restrict__ tab_all,
                                                       const double* __restrict__ aux_all, const double* __restrict__ acc_all, int C, double w_def,
                                                       double* __restrict__ cost, double* __restrict__ grad_all /* may be null */) {
  extern __shared__ __align__(16) unsigned char smem[];
  const ModelC& M = *Mp;
  const int c = blockIdx.x;
  if (c >= C) return;
  epilogue_warp(M, X + (size_t)c * M.nx, tab_all + (size_t)c * M.tab_len, aux_all + (size_t)c * M.aux_len, acc_all + (size_t)c * M.acc_len, w_def,
                cost + c, grad_all ? grad_all + (size_t)c * M.nx : nullptr, reinterpret_cast<double*>(smem), threadIdx.x);
}

// ------------------------------------------------------------------------------------------------
// Point ingestion helpers shared with flash_ingest.cuh (which holds the pipeline: bounds, Hilbert keys, radix sort,
// gather + group balls): fp64 or fp32 [n][3] AoS (the memory of a Julia Vector{SVector{3,Float64}}) -> float4 rounded once.
template <class T>
__device__ __forceinline__ float3 load_xyz(const T* __restrict__ src, long long i) {
  return make_float3((float)src[3 * i], (float)src[3 * i + 1], (float)src[3 * i + 2]);
}
__device__ __forceinline__ unsigned ordered_bits(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ inline float from_ordered_bits(unsigned u) {
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}
// plain conversion of query points (flash_skin_eval)
template <class T>
__global__ void convert_points_kernel(const T* __restrict__ src, float4* __restrict__ dst, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { const float3 p = load_xyz(src, i); dst[i] = make_float4(p.x, p.y, p.z, 0.f); }
}

// ------------------------------------------------------------------------------------------------
// Ray marching (doRaycast, src/depthsensors.jl:56-81): one thread per ray, march state in fp64, the
// field evaluated by the same eval_point as the cost path.  rays are sensor-frame, cam = [R|t] rowmajor.
struct RayParams {
  const HotModel* hot;
  const F4* geo_f4;
  const uint32_t* geo_u32;
  const float* grid_pool;
  const double* tab;      // config 0
  const double* rays;     // [R][3]
  long long R;
  double cam[12];
  double* depths;         // [R], NaN = miss
  int* block_hits;        // [gridDim.x]
  int use_grid;
};

__device__ __forceinline__ double field_at(const HotModel& H, const GeoView& G, const double* tab, const float* tabf, double ox, double oy, double oz,
                                            double rx, double ry, double rz, double dist, unsigned cand, int use_grid) {
  PointEval e;
  eval_point(H, G, tab, tabf, fma(dist, rx, ox), fma(dist, ry, oy), fma(dist, rz, oz), cand, -1, use_grid != 0, false, e);
  return e.is_skin ? e.sk.s : (double)e.d;
}

__global__ void __launch_bounds__(BLOCK, 2) raycast_kernel(const RayParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int tid = threadIdx.x;
  const HotModel* hot_g = p.hot;
  const SmemLayout L = smem_layout(hot_g->tab_len, 0, hot_g->tabf_len, false);
  stage_block(smem, L, hot_g, p.tab, hot_g->tab_len);
  __syncthreads();
  const HotModel& H = *reinterpret_cast<const HotModel*>(smem + L.off_hot);
  GeoView G;
  G.f4 = p.geo_f4;
  G.u32 = p.geo_u32;
  G.grid_pool = p.grid_pool;
  const double* tab = reinterpret_cast<const double*>(smem + L.off_tab);
  const float* tabf = reinterpret_cast<const float*>(smem + L.off_tabf);
  const unsigned cand = H.n_convex >= 32 ? 0xffffffffu : ((1u << H.n_convex) - 1u);
  const long long i = (long long)blockIdx.x * BLOCK + tid;
  bool hit = false;
  if (i < p.R) {
    // rays_in_world (:83-86) then normalize (:93)
    double sx = p.rays[3 * i], sy = p.rays[3 * i + 1], sz = p.rays[3 * i + 2];
    double rx = p.cam[0] * sx + p.cam[1] * sy + p.cam[2] * sz;
    double ry = p.cam[4] * sx + p.cam[5] * sy + p.cam[6] * sz;
    double rz = p.cam[8] * sx + p.cam[9] * sy + p.cam[10] * sz;
    const double inv = 1.0 / sqrt(rx * rx + ry * ry + rz * rz);
    rx *= inv; ry *= inv; rz *= inv;
    const double ox = p.cam[3], oy = p.cam[7], oz = p.cam[11];
    const double EPS = 1e-5, SAFE_RATE = 0.4;
    double dist = 0.0, est = -1.0;
    double last = field_at(H, G, tab, tabf, ox, oy, oz, rx, ry, rz, dist, cand, p.use_grid);
    int k = 0;
    while (fabs(last) > EPS && k < 60) {
      double step = -last / est;
      const double sg = step > 0 ? 1.0 : (step < 0 ? -1.0 : 0.0);
      step = sg * fmin(SAFE_RATE, fabs(step));
      dist += step;
      const double v = field_at(H, G, tab, tabf, ox, oy, oz, rx, ry, rz, dist, cand, p.use_grid);
      est = (v - last) / step;
      last = v;
      ++k;
    }
    // the reference re-evaluates the field at the final sample; `last` is that value
    hit = !(fabs(last) > 1000 * EPS);
    p.depths[i] = hit ? dist : __longlong_as_double(0x7ff8000000000000LL);
  }
  const int cnt = __syncthreads_count(hit);
  if (tid == 0) p.block_hits[blockIdx.x] = cnt;
}

// exclusive scan of per-block hit counts (single block; nblocks up to a few thousand)
__global__ void scan_blocks_kernel(int* __restrict__ counts, int n, long long* __restrict__ total) {
  __shared__ int carry;
  __shared__ int wsum[32];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n; base += blockDim.x) {
    const int i = base + threadIdx.x;
    int v = i < n ? counts[i] : 0;
    int x = v;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    if (w == 0) {
      int s = lane < (int)(blockDim.x >> 5) ? wsum[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
      wsum[lane] = s;
    }
    __syncthreads();
    const int prev = (w > 0 ? wsum[w - 1] : 0) + carry;
    if (i < n) counts[i] = prev + x - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = prev + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}

// hits in ray order: sensor_tform(dist * normalize(sensor.rays[row, col])), src/depthsensors.jl:108
__global__ void __launch_bounds__(BLOCK) compact_hits_kernel(const double* __restrict__ rays, const double* __restrict__ depths, long long R,
                                                             const int* __restrict__ block_base, double* __restrict__ hits, float4* __restrict__ hits_f4,
                                                             double c0, double c1, double c2, double c3, double c4, double c5, double c6, double c7,
                                                             double c8, double c9, double c10, double c11) {
  __shared__ int wbase[WARPS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long i = (long long)blockIdx.x * BLOCK + tid;
  double d = i < R ? depths[i] : __longlong_as_double(0x7ff8000000000000LL);
  const bool hit = d == d;
  const unsigned m = __ballot_sync(0xffffffffu, hit);
  if (lane == 0) wbase[warp] = __popc(m);
  __syncthreads();
  int base = block_base[blockIdx.x];
  for (int w = 0; w < warp; ++w) base += wbase[w];
  if (hit) {
    const long long k = base + __popc(m & ((1u << lane) - 1u));
    double sx = rays[3 * i], sy = rays[3 * i + 1], sz = rays[3 * i + 2];
    const double s = d / sqrt(sx * sx + sy * sy + sz * sz);
    sx *= s; sy *= s; sz *= s;
    const double wx = c0 * sx + c1 * sy + c2 * sz + c3, wy = c4 * sx + c5 * sy + c6 * sz + c7, wz = c8 * sx + c9 * sy + c10 * sz + c11;
    if (hits) { hits[3 * k] = wx; hits[3 * k + 1] = wy; hits[3 * k + 2] = wz; }
    if (hits_f4) hits_f4[k] = make_float4((float)wx, (float)wy, (float)wz, 0.f);
  }
}

// One NaiveSolver update on one warp (flash_estimate_state in flashsdf.h): step = -rate * (g / N) ./ divisors, its norm
// clamped to max_step; returns true (and leaves x alone) when |g / N| < tol.
__device__ __forceinline__ bool naive_step_warp(int nx, const double* g, double* x, double inv_n, double rate, double max_step, double tol,
                                                const double* divisors, int lane) {
  double gn = 0, sn = 0;
  double stv[8];                       // this lane's step components (nx <= 256): divided once, applied below
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int k = lane + 32 * q;
    stv[q] = 0.0;
    if (k < nx) {
      const double gk = g[k] * inv_n;
      const double st = -rate * gk / (divisors ? divisors[k] : 1.0);
      stv[q] = st;
      gn += gk * gk; sn += st * st;
    }
  }
  for (int k = lane + 256; k < nx; k += 32) {   // states beyond the register window (up to 64 + 3 * 4 * 64)
    const double gk = g[k] * inv_n;
    const double st = -rate * gk / (divisors ? divisors[k] : 1.0);
    gn += gk * gk; sn += st * st;
  }
  gn = warp_sum(gn); sn = sqrt(warp_sum(sn));
  if (tol > 0 && sqrt(gn) < tol) return true;
  const double sc = sn > max_step ? max_step / sn : 1.0;
#pragma unroll
  for (int q = 0; q < 8; ++q) { const int k = lane + 32 * q; if (k < nx) x[k] += sc * stv[q]; }
  for (int k = lane + 256; k < nx; k += 32) x[k] += sc * (-rate * g[k] * inv_n / (divisors ? divisors[k] : 1.0));
  return false;
}

// ------------------------------------------------------------------------------------------------
// Tracking iteration tail in ONE launch, one block per config: epilogue (cost + gradient of the pass the cost
// kernel just finished; warp 0) -> NaiveSolver step (flash_estimate_state in flashsdf.h; warp 0) -> prepare of
// the NEXT pass (FK, RBF system, zeroed accumulators; all PREP_T threads).  A gradient-descent iteration is then
// 2 launches (cost_kernel, this) instead of 4; the three single-warp kernels it replaces were 51 of the 68 us
// of an iteration on K3.
__global__ void __launch_bounds__(PREP_T) iterate_kernel(const ModelC* Mp, double* X, double* tab_all, double* aux_all, double* acc_all, int C, double w_def,
                                                          double* cost, double* grad_all, double inv_n, double rate, double max_step, double tol,
                                                          const double* divisors, int* done, int* iters, int* status) {
  extern __shared__ __align__(16) unsigned char smem[];
  double* sm = reinterpret_cast<double*>(smem);
  const ModelC& M = *Mp;
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
  if (c >= C) return;
  double* x = X + (size_t)c * M.nx;
  double* tab = tab_all + (size_t)c * M.tab_len;
  double* aux = aux_all + (size_t)c * M.aux_len;
  double* acc = acc_all + (size_t)c * M.acc_len;
  double* g = grad_all + (size_t)c * M.nx;
#ifdef FLASH_ITER_TIMING
  long long t0 = clock64(), t1 = 0, t2 = 0;
#endif
  epilogue_preload(M, tab, aux, sm, tid, PREP_T);   // 320 threads fetch the first skin's inverse while nothing else runs
  __syncthreads();
  if (tid < 32) {
    epilogue_warp(M, x, tab, aux, acc, w_def, cost + c, g, sm, lane, true);
    __syncwarp();
#ifdef FLASH_ITER_TIMING
    t1 = clock64();
#endif
    if (!done[c]) {
      if (naive_step_warp(M.nx, g, x, inv_n, rate, max_step, tol, divisors, lane)) { if (lane == 0) done[c] = 1; }
      else if (lane == 0) iters[c] += 1;
    }
  }
#ifdef FLASH_ITER_TIMING
  t2 = clock64();
#endif
  __syncthreads();   // x, and the end of every read of tab / aux / acc, are visible to the whole block
  prepare_block(M, x, tab, aux, acc, status, sm, tid, PREP_T);
#ifdef FLASH_ITER_TIMING
  if (tid == 0 && c == 0 && iters[0] >= 20 && iters[0] < 23) printf("iterate_kernel cycles: epilogue %lld step %lld prepare %lld\n", t1 - t0, t2 - t1, clock64() - t2);
#endif
}

// ------------------------------------------------------------------------------------------------
// The whole solver loop of one frame in ONE launch (flash_estimate_state: one configuration, a model whose point loop
// needs no group records, one device).  A persistent grid of at most one block per SM, all co-resident (cooperative
// launch).  Per iteration:
//   cost pass   every block runs cost_groups over its own tile out of tables in ITS shared memory and adds the sums of its
//               warps to the pass's totals (fp64 atomics; deterministic mode: stores them per block)
//   barrier     one grid-wide barrier (arrive counter + spin; bounded, so a lost block ends the call instead of hanging)
//   tail        EVERY block reads the totals (deterministic mode: adds the blocks' sums in a fixed order), then runs the epilogue, the NaiveSolver
//               step and the prepare of the next pass redundantly on its own copy of the state: the other SMs would only
//               wait, and what the next pass reads (centres, weights, inverse) is already where it is needed.
// The two-kernel loop it replaces (cost_kernel, iterate_kernel per iteration) spent ~9 us per iteration on launches and on
// moving tables, inverse and sums through global memory.  Same arithmetic in the same order per block
// as the two-kernel loop in deterministic mode, except that the sums over the points are grouped by tile.
struct TrackParams {
  CostParams cp;            // hot, geometry, pts, n_pts, pts_per_block, use_grid (tab / acc / rec / part unused)
  const ModelC* model;
  double* X;                // [nx]  in: start, out: estimate
  double* part;             // deterministic: [2][acc_len][gridDim.x] block sums; else [3][acc_len] totals, zero at launch
  unsigned* bar;            // grid barrier counter, zero at launch
  double* cost;             // [1] cost at the final state (want_cost)
  int* iters; int* done; int* status;
  const double* divisors;   // may be null
  double w_def, inv_n, rate, max_step, tol;
  int iteration_limit, want_cost;
};
struct TrackLayout { SmemLayout L; size_t off_x, off_aux, off_grad, off_sum, off_scr, total; };
__host__ __device__ inline TrackLayout track_layout(const ModelC& M) {
  TrackLayout T;
  T.L = smem_layout(M.tab_len, M.acc_len, M.tabf_len, M.n_skins > 0, CWARPS);
  size_t o = T.L.total;
  T.off_x = o; o = align16(o + sizeof(double) * (size_t)M.nx);
  T.off_aux = o; o = align16(o + sizeof(double) * (size_t)(M.aux_len > 0 ? M.aux_len : 1));
  T.off_grad = o; o = align16(o + sizeof(double) * (size_t)M.nx);
  T.off_sum = o; o = align16(o + sizeof(double) * (size_t)M.acc_len);
  const size_t a = prepare_smem_doubles(M), b = epilogue_smem_doubles(M);
  T.off_scr = o; o = align16(o + sizeof(double) * (a > b ? a : b));
  T.total = o;
  return T;
}
constexpr unsigned TRACK_SPIN_LIMIT = 1u << 22;   // seconds of polling: far beyond any pass; a grid that is not co-resident gives up
__device__ __forceinline__ bool track_grid_barrier(unsigned* bar, unsigned target, int* s_lost) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    unsigned spins = 0;
    while (*(volatile unsigned*)bar < target) { if (++spins > TRACK_SPIN_LIMIT) { *s_lost = 1; break; } __nanosleep(32); }
    __threadfence();
  }
  __syncthreads();
  return *s_lost == 0;
}
// DET: deterministic mode (static schedule, block sums gathered in a fixed order).  STATIC: static schedule only -- chosen
// when a block has at most one group per warp: warps 0 .. g-1 then sit evenly on the four schedulers (9 groups: 3+2+2+2),
// while the first g warps to win the hand-out race may not (the pass is bound by the fp64 pipe of the fullest scheduler).
template <bool DET, bool STATIC>
__global__ void __launch_bounds__(CBLOCK, 1) track_kernel(const TrackParams t) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const ModelC& M = *t.model;
  const TrackLayout T = track_layout(M);
  const SmemLayout& L = T.L;
  const int acc_len = M.acc_len, nx = M.nx, nblk = gridDim.x;
  __shared__ int s_next, s_lost, s_done;
  double* s_x = reinterpret_cast<double*>(smem + T.off_x);
  double* s_aux = reinterpret_cast<double*>(smem + T.off_aux);
  double* s_grad = reinterpret_cast<double*>(smem + T.off_grad);
  double* s_sum = reinterpret_cast<double*>(smem + T.off_sum);
  double* s_scr = reinterpret_cast<double*>(smem + T.off_scr);
  double* tab = reinterpret_cast<double*>(smem + L.off_tab);
  float* tabf = reinterpret_cast<float*>(smem + L.off_tabf);
  double* acc_all = reinterpret_cast<double*>(smem + L.off_acc);
  double* acc_w = acc_all + (size_t)warp * acc_len;
  double* queue = reinterpret_cast<double*>(smem + L.off_queue) + (size_t)warp * 32 * QENT;
  int* qslot = reinterpret_cast<int*>(smem + L.off_qslot) + warp * 32;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(t.cp.hot);
    uint32_t* dst = reinterpret_cast<uint32_t*>(smem + L.off_hot);
    for (int i = tid; i < (int)(sizeof(HotModel) / 4); i += CBLOCK) dst[i] = src[i];
    for (int i = tid; i < nx; i += CBLOCK) s_x[i] = t.X[i];
    if (tid == 0) { s_lost = 0; s_done = 0; }
  }
  __syncthreads();
  const HotModel& H = *reinterpret_cast<const HotModel*>(smem + L.off_hot);
  GeoView G;
  G.f4 = t.cp.geo_f4; G.u32 = t.cp.geo_u32; G.grid_pool = t.cp.grid_pool;
  const long long n = t.cp.n_pts;
  const long long tile0 = (long long)blockIdx.x * t.cp.pts_per_block;
  long long tile1 = tile0 + t.cp.pts_per_block;
  if (tile1 > n) tile1 = n;
  int iters = 0;
  unsigned passes = 0;
  const int last = t.iteration_limit + (t.want_cost ? 1 : 0);   // the extra pass is the cost at the final state
#ifdef FLASH_TRACK_TIMING
  long long tk[8];
#define FLASH_TK(i) tk[i] = clock64()
#else
#define FLASH_TK(i) ((void)0)
#endif
  for (int it = 0; it < last; ++it) {
    const bool final_pass = it == t.iteration_limit;
    FLASH_TK(0);
    // ---- tables of this pass from this block's copy of the state
    prepare_block(M, s_x, tab, s_aux, nullptr, t.status, s_scr, tid, CBLOCK);
    __syncthreads();
    tabf_fill(H, tab, tabf, tid, CBLOCK);
    for (int i = tid; i < acc_len * CWARPS; i += CBLOCK) acc_all[i] = 0.0;
    if (tid == 0) s_next = 0;
    __syncthreads();
    FLASH_TK(1);
    // ---- cost pass over this block's tile
    double cost_local = 0.0;
    if (tile0 < tile1) cost_local = cost_groups<true, false, false, DET || STATIC, 1, true>(t.cp, H, G, tab, tabf, acc_w, queue, qslot, &s_next, 0, H.n_convex, 0ll, tile0, tile1, lane, warp);
    cost_local = warp_sum(cost_local);
    if (lane == 0) acc_w[0] += cost_local;
#ifdef FLASH_TRACK_TIMING
    __shared__ long long s_wclk[CWARPS];
    if (lane == 0) s_wclk[warp] = clock64();
#endif
    __syncthreads();
    FLASH_TK(2);
#ifdef FLASH_TRACK_TIMING
    if (tid == 0) { long long mx = 0; for (int w = 0; w < CWARPS; ++w) mx = s_wclk[w] > mx ? s_wclk[w] : mx; tk[7] = mx - tk[1]; }
#endif
    if (DET) {      // this block's sums, layout [pass & 1][accumulator][block]: the gather below reads whole rows
      double* mine = t.part + (size_t)(passes & 1u) * nblk * acc_len + blockIdx.x;
      for (int l = tid; l < acc_len; l += CBLOCK) {
        double s = 0;
#pragma unroll
        for (int w = 0; w < CWARPS; ++w) s += acc_all[(size_t)w * acc_len + l];
        __stcg(mine + (size_t)l * nblk, s);
      }
    } else {        // fp64 atomics into one of three rotating totals (zeroed two barriers ahead of their next use)
      double* tot = t.part + (size_t)(passes % 3u) * acc_len;
      for (int l = tid; l < acc_len; l += CBLOCK) {
        double s = 0;
#pragma unroll
        for (int w = 0; w < CWARPS; ++w) s += acc_all[(size_t)w * acc_len + l];
        if (s != 0.0) atomicAdd(tot + l, s);
      }
    }
    ++passes;
    if (!track_grid_barrier(t.bar, passes * (unsigned)nblk, &s_lost)) break;
    FLASH_TK(3);
    if (DET) {      // sums over the blocks in a fixed order: warp w takes accumulators w, w + 16, ..., four at a time
      const double* all = t.part + (size_t)((passes - 1u) & 1u) * nblk * acc_len;
      for (int l0 = 0; l0 < acc_len; l0 += 4 * CWARPS) {
        double s[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int l = l0 + warp + k * CWARPS;
          if (l < acc_len) for (int b = lane; b < nblk; b += 32) s[k] += __ldcg(all + (size_t)l * nblk + b);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int l = l0 + warp + k * CWARPS;
          const double v = warp_sum(s[k]);
          if (l < acc_len && lane == 0) s_sum[l] = v;
        }
      }
    } else {
      const double* tot = t.part + (size_t)((passes - 1u) % 3u) * acc_len;
      for (int l = tid; l < acc_len; l += CBLOCK) s_sum[l] = __ldcg(tot + l);
      if (blockIdx.x == 0) { double* z = t.part + (size_t)((passes + 1u) % 3u) * acc_len; for (int l = tid; l < acc_len; l += CBLOCK) z[l] = 0.0; }
    }
    __syncthreads();
    FLASH_TK(4);
    // ---- epilogue and solver step on warp 0
    if (tid < 32) {
      double cst = 0.0;
      epilogue_warp(M, s_x, tab, s_aux, s_sum, t.w_def, &cst, final_pass ? nullptr : s_grad, s_scr, lane, false, true);
      __syncwarp();
      if (final_pass) { if (blockIdx.x == 0 && lane == 0) t.cost[0] = cst; }
      else if (naive_step_warp(nx, s_grad, s_x, t.inv_n, t.rate, t.max_step, t.tol, t.divisors, lane)) {   // converged: x stays, and cst is the cost there
        if (lane == 0) { s_done = 1; if (blockIdx.x == 0 && t.want_cost) t.cost[0] = cst; }
      } else ++iters;
    }
    __syncthreads();
#ifdef FLASH_TRACK_TIMING
    if (tid == 0 && it == 20)
      printf("track_kernel block %d cycles: prepare %lld cost %lld (slowest warp %lld) sums+barrier %lld gather %lld epilogue+step %lld\n", (int)blockIdx.x, tk[1] - tk[0], tk[2] - tk[1], tk[7], tk[3] - tk[2], tk[4] - tk[3], clock64() - tk[4]);
#endif
    if (final_pass || s_done) break;
  }
  if (blockIdx.x == 0) {
    for (int i = tid; i < nx; i += CBLOCK) t.X[i] = s_x[i];
    if (tid == 0) { t.iters[0] += iters; t.done[0] = s_done; if (s_lost) atomicExch(t.status, 2); }
  } else if (tid == 0 && s_lost) atomicExch(t.status, 2);
}

// ------------------------------------------------------------------------------------------------
// grid mode: bake the exact hull distance at every node of a geometry-frame grid
__global__ void bake_grid_kernel(const HotModel* __restrict__ hot, const F4* __restrict__ geo_f4, const uint32_t* __restrict__ geo_u32, int j, float* __restrict__ pool) {
  const ConvexC& C = hot->cvx[j];
  const long long n = (long long)C.g_dim[0] * C.g_dim[1] * C.g_dim[2];
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int ix = (int)(i % C.g_dim[0]), iy = (int)((i / C.g_dim[0]) % C.g_dim[1]), iz = (int)(i / ((long long)C.g_dim[0] * C.g_dim[1]));
  const float h = 1.0f / C.g_inv_h;
  F3 y = f3(C.g_org[0] + ix * h, C.g_org[1] + iy * h, C.g_org[2] + iz * h);
  F3 n3;
  GeoView G; G.f4 = geo_f4; G.u32 = geo_u32; G.grid_pool = nullptr;
  pool[C.g_off + i] = hull_sd(hull_view(G, C), y, n3);
}
// max |trilinear - exact| over cell centres (the worst place for trilinear interpolation)
__global__ void grid_error_kernel(const HotModel* __restrict__ hot, const F4* __restrict__ geo_f4, const uint32_t* __restrict__ geo_u32, int j,
                                  const float* __restrict__ pool, unsigned int* __restrict__ max_err_bits) {
  const ConvexC& C = hot->cvx[j];
  const int cx = C.g_dim[0] - 1, cy = C.g_dim[1] - 1, cz = C.g_dim[2] - 1;
  const long long n = (long long)cx * cy * cz;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float err = 0.f;
  if (i < n) {
    const int ix = (int)(i % cx), iy = (int)((i / cx) % cy), iz = (int)(i / ((long long)cx * cy));
    const float h = 1.0f / C.g_inv_h;
    F3 y = f3(C.g_org[0] + (ix + 0.5f) * h, C.g_org[1] + (iy + 0.5f) * h, C.g_org[2] + (iz + 0.5f) * h);
    F3 n3, n4; float dg;
    GeoView G; G.f4 = geo_f4; G.u32 = geo_u32; G.grid_pool = nullptr;
    float de = hull_sd(hull_view(G, C), y, n3);
    if (grid_sd(C, pool, y, dg, n4)) err = fabsf(dg - de);
  }
  err = fmaxf(err, __shfl_xor_sync(0xffffffffu, err, 16));
  err = fmaxf(err, __shfl_xor_sync(0xffffffffu, err, 8));
  err = fmaxf(err, __shfl_xor_sync(0xffffffffu, err, 4));
  err = fmaxf(err, __shfl_xor_sync(0xffffffffu, err, 2));
  err = fmaxf(err, __shfl_xor_sync(0xffffffffu, err, 1));
  if ((threadIdx.x & 31) == 0 && err > 0.f) atomicMax(max_err_bits, __float_as_uint(err));
}

// ------------------------------------------------------------------------------------------------
// Headless iso-surface extraction (row F4): what `draw` obtains from contour_mesh(surface, lb, ub, 0.0, ..)
// (src/Flash.jl:318-326), written to a triangle buffer instead of a viewer.
//   field_grid_kernel  samples one surface's field (or the min over all surfaces, surface < 0) at the nodes
//                      of a regular world-frame grid with the SAME evaluators as the cost path
//   march_tets_kernel  marching tetrahedra (6 per cell, no case table): 0-2 triangles per tetrahedron,
//                      vertices by linear interpolation along the edges, oriented towards increasing field
struct GridParams {
  const HotModel* hot;
  const F4* geo_f4;
  const uint32_t* geo_u32;
  const double* tab;      // config 0
  double org[3]; double h; int dim[3];
  int surface;            // surface index, or < 0 for skin(state) = min over all
  float* field;           // [dim2][dim1][dim0]
};
__global__ void __launch_bounds__(BLOCK, 2) field_grid_kernel(const GridParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const HotModel* hot_g = p.hot;
  const SmemLayout L = smem_layout(hot_g->tab_len, 0, hot_g->tabf_len, false);
  stage_block(smem, L, hot_g, p.tab, hot_g->tab_len);
  __syncthreads();
  const HotModel& H = *reinterpret_cast<const HotModel*>(smem + L.off_hot);
  GeoView G;
  G.f4 = p.geo_f4;
  G.u32 = p.geo_u32;
  G.grid_pool = nullptr;
  const double* tab = reinterpret_cast<const double*>(smem + L.off_tab);
  const float* tabf = reinterpret_cast<const float*>(smem + L.off_tabf);
  const long long n = (long long)p.dim[0] * p.dim[1] * p.dim[2];
  const long long i = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (i >= n) return;
  const int ix = (int)(i % p.dim[0]), iy = (int)((i / p.dim[0]) % p.dim[1]), iz = (int)(i / ((long long)p.dim[0] * p.dim[1]));
  const double px = p.org[0] + ix * p.h, py = p.org[1] + iy * p.h, pz = p.org[2] + iz * p.h;
  double v;
  if (p.surface < 0) {
    PointEval e;
    const unsigned cand = H.n_convex >= 32 ? 0xffffffffu : ((1u << H.n_convex) - 1u);
    eval_point(H, G, tab, tabf, px, py, pz, cand, -1, false, false, e);
    v = e.is_skin ? e.sk.s : (double)e.d;
  } else {
    v = 3.0e38;
    for (int k = 0; k < H.n_skins; ++k)
      if (H.skin[k].surface == p.surface) {
        const double* c = tab + H.skin[k].tab_off;
        v = skin_eval(c, c + 3 * H.skin[k].npts, H.skin[k].npts, H.normalize != 0, px, py, pz).s;
      }
    for (int j = 0; j < H.n_convex; ++j)
      if (H.cvx[j].surface == p.surface) { F3 nn; v = (double)hull_sd(hull_view(G, H.cvx[j]), to_geometry(tab + 12 * j, px, py, pz), nn); }
  }
  p.field[i] = (float)v;
}

struct MarchParams {
  const float* field; double org[3]; double h; int dim[3]; float iso;
  float* tris;                  // [cap][9] world-frame vertices, may be null (count only)
  unsigned long long cap;
  unsigned long long* count;    // triangles produced (may exceed cap: the caller re-runs with a larger buffer)
};
__device__ __forceinline__ void march_emit(const MarchParams& p, const float a[3], const float b[3], const float c[3], const float towards[3]) {
  // orient: normal . (direction of increasing field) > 0
  const float e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
  const float nx = e1[1] * e2[2] - e1[2] * e2[1], ny = e1[2] * e2[0] - e1[0] * e2[2], nz = e1[0] * e2[1] - e1[1] * e2[0];
  if (nx == 0.f && ny == 0.f && nz == 0.f) return;   // degenerate (iso value exactly on a node)
  const bool flip = nx * towards[0] + ny * towards[1] + nz * towards[2] < 0.f;
  const unsigned long long k = atomicAdd(p.count, 1ull);
  if (p.tris == nullptr || k >= p.cap) return;
  float* o = p.tris + 9 * k;
  const float* v1 = flip ? c : b; const float* v2 = flip ? b : c;
  o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; o[3] = v1[0]; o[4] = v1[1]; o[5] = v1[2]; o[6] = v2[0]; o[7] = v2[1]; o[8] = v2[2];
}
__global__ void __launch_bounds__(256) march_tets_kernel(const MarchParams p) {
  const int cx = p.dim[0] - 1, cy = p.dim[1] - 1, cz = p.dim[2] - 1;
  const long long n = (long long)cx * cy * cz;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int ix = (int)(i % cx), iy = (int)((i / cx) % cy), iz = (int)(i / ((long long)cx * cy));
  float val[8], pos[8][3];
  bool any_in = false, any_out = false;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int dx = k & 1, dy = (k >> 1) & 1, dz = k >> 2;
    val[k] = p.field[((long long)(iz + dz) * p.dim[1] + (iy + dy)) * p.dim[0] + (ix + dx)] - p.iso;
    pos[k][0] = (float)(p.org[0] + (ix + dx) * p.h); pos[k][1] = (float)(p.org[1] + (iy + dy) * p.h); pos[k][2] = (float)(p.org[2] + (iz + dz) * p.h);
    any_in |= val[k] < 0.f; any_out |= !(val[k] < 0.f);
  }
  if (!(any_in && any_out)) return;
  // six tetrahedra around the main diagonal 0-7
  const int tets[6][4] = {{0, 1, 3, 7}, {0, 3, 2, 7}, {0, 2, 6, 7}, {0, 6, 4, 7}, {0, 4, 5, 7}, {0, 5, 1, 7}};
  for (int t = 0; t < 6; ++t) {
    int in[4], out[4], ni = 0, no = 0;
    for (int k = 0; k < 4; ++k) { const int v = tets[t][k]; if (val[v] < 0.f) in[ni++] = v; else out[no++] = v; }
    if (ni == 0 || ni == 4) continue;
    auto cut = [&](int a, int b, float* o) {   // a inside, b outside
      const float w = val[a] / (val[a] - val[b]);
      o[0] = pos[a][0] + w * (pos[b][0] - pos[a][0]); o[1] = pos[a][1] + w * (pos[b][1] - pos[a][1]); o[2] = pos[a][2] + w * (pos[b][2] - pos[a][2]);
    };
    float dir[3] = {0.f, 0.f, 0.f};            // mean outside corner - mean inside corner
    for (int k = 0; k < no; ++k) for (int a = 0; a < 3; ++a) dir[a] += pos[out[k]][a] / no;
    for (int k = 0; k < ni; ++k) for (int a = 0; a < 3; ++a) dir[a] -= pos[in[k]][a] / ni;
    float q0[3], q1[3], q2[3], q3[3];
    if (ni == 1) { cut(in[0], out[0], q0); cut(in[0], out[1], q1); cut(in[0], out[2], q2); march_emit(p, q0, q1, q2, dir); }
    else if (ni == 3) { cut(in[0], out[0], q0); cut(in[1], out[0], q1); cut(in[2], out[0], q2); march_emit(p, q0, q1, q2, dir); }
    else {  // 2-2: a quad
      cut(in[0], out[0], q0); cut(in[0], out[1], q1); cut(in[1], out[1], q2); cut(in[1], out[0], q3);
      march_emit(p, q0, q1, q2, dir); march_emit(p, q0, q2, q3, dir);
    }
  }
}

}  // namespace flashk
