// flashsdf.cu -- C ABI of libflashsdf (include/flashsdf.h): context, device memory, streams,
// kernel launches, NCCL all-reduce.  There is no CPU compute path in this file or anywhere in the
// library: without a CUDA device the compute entry points return FLASH_ERR_NO_DEVICE.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/flashsdf.h"
#include "flash_host.h"
#include "flash_urdf.h"
#include "flash_kernels.cuh"
#include "flash_ingest.cuh"

using namespace flashdev;
using namespace flashhost;
using namespace flashk;

// ------------------------------------------------------------------------------------------------
// NCCL is resolved at run time (dlopen) so that single-GPU users -- the Julia host -- need no NCCL.
namespace {
typedef struct { char internal[128]; } nccl_uid;
typedef void* nccl_comm;
struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(nccl_uid*) = nullptr;
  int (*CommInitRank)(nccl_comm*, int, nccl_uid, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm, cudaStream_t) = nullptr;
  int (*CommDestroy)(nccl_comm) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
bool load_nccl(std::string& err) {
  if (g_nccl.handle) return true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
  if (!h) { err = std::string("cannot dlopen libnccl.so.2: ") + dlerror(); return false; }
  NcclApi a; a.handle = h;
  a.GetUniqueId = (int (*)(nccl_uid*))dlsym(h, "ncclGetUniqueId");
  a.CommInitRank = (int (*)(nccl_comm*, int, nccl_uid, int))dlsym(h, "ncclCommInitRank");
  a.AllReduce = (int (*)(const void*, void*, size_t, int, int, nccl_comm, cudaStream_t))dlsym(h, "ncclAllReduce");
  a.CommDestroy = (int (*)(nccl_comm))dlsym(h, "ncclCommDestroy");
  a.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
  if (!a.GetUniqueId || !a.CommInitRank || !a.AllReduce || !a.CommDestroy) { err = "libnccl.so.2 lacks required symbols"; return false; }
  g_nccl = a;
  return true;
}
constexpr int NCCL_FLOAT64 = 8, NCCL_SUM = 0;
thread_local std::string g_create_err;
}  // namespace

struct flash_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  HostModel hm;
  HotModel hot;
  ModelC* d_model = nullptr;
  HotModel* d_hot = nullptr;
  F4* d_geo_f4 = nullptr;
  uint32_t* d_geo_u32 = nullptr;
  float* d_grid_pool = nullptr;
  size_t grid_pool_floats = 0;
  std::vector<double> grid_err;  // per convex slot
  bool use_grid = false;
  // points: two sets (front = the frame(s) being evaluated, back = the one being uploaded / sorted on the copy stream)
  struct PointSet {
    float4* d_pts = nullptr; long long cap_pts = 0; bool external = false;
    int* d_perm = nullptr; size_t cap_perm = 0; bool perm_on = false;
    float4* d_grp = nullptr; size_t cap_grp = 0; bool grp_on = false;
    long long* d_frames = nullptr;     // device: off[0 .. MAX_FRAMES], cnt[MAX_FRAMES + 1 ..]
    FrameTable ft;                     // the same on the host (+ offsets in the caller-order staging buffer)
    long long n = 0, max_cnt = 0;      // points of all frames; of the largest frame
    bool set = false;
    cudaEvent_t ready = nullptr;       // copy stream: upload + sort + group balls finished
    cudaEvent_t retired = nullptr;     // compute stream: every kernel that read this set has been issued before this event
    bool retired_valid = false;
  };
  PointSet ps[2];
  int cur = 0;
  bool back_pending = false;           // flash_points_set_async has filled the back set, flash_points_swap not yet called
  cudaStream_t copy_stream = nullptr;
  void* d_stage = nullptr; size_t cap_stage = 0;           // caller-order [n][3] as uploaded
  void* h_ring[2] = {nullptr, nullptr}; size_t cap_ring[2] = {0, 0}; cudaEvent_t ring_done[2] = {nullptr, nullptr}; int ring_next = 0;   // pinned staging for pageable callers
  unsigned* d_keys[2] = {nullptr, nullptr}; int* d_idx[2] = {nullptr, nullptr}; size_t cap_sort = 0;
  int* d_table = nullptr; size_t cap_table = 0; int* d_sums = nullptr; unsigned* d_bounds = nullptr;   // [6 * MAX_FRAMES] bounds + [1] scan counter
  bool sort_points = true; bool group_cull = true;
  // per-batch state
  int cap_C = 0;
  double *d_X = nullptr, *d_tab = nullptr, *d_aux = nullptr, *d_acc = nullptr, *d_cost = nullptr, *d_grad = nullptr, *d_div = nullptr;
  int *d_status = nullptr, *d_done = nullptr, *d_iters = nullptr;
  double* h_pin = nullptr; size_t cap_pin = 0;  // pinned staging for X / cost / grad
  int32_t* d_argmin = nullptr; size_t cap_argmin = 0;
  uint4* d_rec = nullptr; size_t cap_rec = 0;   // group records of group_kernel, [C][rec_stride]
  double* d_part = nullptr; size_t cap_part = 0; bool deterministic = false;   // per-tile partial sums [C][tiles][acc_len]
  unsigned* d_bar = nullptr;   // grid barrier counter of track_kernel
  int* d_order = nullptr; size_t cap_order = 0; unsigned* d_tile_w = nullptr;   // block row -> tile of cost_kernel (tile_order_kernel), work estimates
  bool exact_groups = true;   // group_kernel evaluates the candidate hulls exactly at the ball centre (flash_dev.h group_prepass)
  // query scratch (skin_eval / raycast)
  double* d_qd = nullptr; size_t cap_qd = 0;
  int* d_qi = nullptr; size_t cap_qi = 0;
  // comm
  nccl_comm comm = nullptr; int rank = 0, world = 1;
  double n_global = -1.0;   // points of all ranks in the current set (flash_estimate_state divides by it); < 0: not reduced yet
  // single-process multi-GPU context (flash_options.n_devices > 1): the parent owns one full sub-context per device and no
  // device state of its own; subs[0] additionally owns the table of every sub's accumulator pointer for the peer reduction
  std::vector<flash_ctx*> subs;
  flash_ctx* parent = nullptr;
  double** d_peer_acc = nullptr; int peer_acc_C = 0;
  cudaEvent_t ev_cost = nullptr, ev_red = nullptr;   // sub: cost kernel done; subs[0]: peer reduction done
  std::vector<std::vector<long long>> seg_off, seg_len;   // parent: per sub, the chunks of the caller's cloud it holds (first point, length)
  long long n_points = 0; bool have_points = false;       // parent: size of the current cloud (false: a batch of frames)
  std::vector<int> frame_sub, frame_slot;            // parent: owner and local index of every frame of the current batch
  // measurement
  bool profiling = false;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // [5]: between group_kernel and cost_kernel
  flash_timing last = {0, 0, 0, 0, 0, 0, 0};
  long long launches = 0;
  int sm_count = 0, clock_khz = 0; long long l2_bytes = 0; int smem_optin = 0;
  std::string err;
};

namespace {

#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                               \
      return FLASH_ERR_CUDA;                                                                       \
    }                                                                                              \
  } while (0)

int fail(flash_ctx* ctx, int code, const std::string& msg) {
  if (ctx) ctx->err = msg; else g_create_err = msg;
  return code;
}

template <class T>
int ensure(flash_ctx* ctx, T*& p, size_t& cap, size_t need) {
  if (need <= cap) return 0;
  if (p) CU(cudaFree(p));
  p = nullptr; cap = 0;
  size_t want = need + need / 8;
  CU(cudaMalloc((void**)&p, want * sizeof(T)));
  cap = want;
  return 0;
}

int ensure_batch(flash_ctx* ctx, int C) {
  if (C <= ctx->cap_C) return 0;
  const ModelC& M = ctx->hm.M;
  double** ptrs[] = {&ctx->d_X, &ctx->d_tab, &ctx->d_aux, &ctx->d_acc, &ctx->d_cost, &ctx->d_grad};
  for (auto pp : ptrs) if (*pp) { CU(cudaFree(*pp)); *pp = nullptr; }
  if (ctx->d_done) { CU(cudaFree(ctx->d_done)); ctx->d_done = nullptr; }
  if (ctx->d_iters) { CU(cudaFree(ctx->d_iters)); ctx->d_iters = nullptr; }
  if (ctx->h_pin) { CU(cudaFreeHost(ctx->h_pin)); ctx->h_pin = nullptr; }
  ctx->cap_C = 0;
  const size_t c = (size_t)C;
  CU(cudaMalloc((void**)&ctx->d_X, c * M.nx * sizeof(double)));
  CU(cudaMalloc((void**)&ctx->d_tab, c * std::max(M.tab_len, 1) * sizeof(double)));
  CU(cudaMalloc((void**)&ctx->d_aux, c * std::max(M.aux_len, 1) * sizeof(double)));
  CU(cudaMalloc((void**)&ctx->d_acc, c * M.acc_len * sizeof(double)));
  CU(cudaMalloc((void**)&ctx->d_cost, c * sizeof(double)));
  CU(cudaMalloc((void**)&ctx->d_grad, c * M.nx * sizeof(double)));
  CU(cudaMalloc((void**)&ctx->d_done, c * sizeof(int)));
  CU(cudaMalloc((void**)&ctx->d_iters, c * sizeof(int)));
  ctx->cap_pin = c * (2 * (size_t)M.nx + 2) + (size_t)M.nx;   // X, cost, grad (+ the solver's divisors)
  CU(cudaHostAlloc((void**)&ctx->h_pin, ctx->cap_pin * sizeof(double), cudaHostAllocDefault));
  ctx->cap_C = C;
  return 0;
}

// Small host -> device uploads (the candidate states, the solver's divisors) are PULLED by a kernel from the ctx's
// pinned staging buffer (mapped into the device's address space by UVA) instead of going through the copy engine:
// a cudaMemcpyAsync would queue behind a frame upload in flight on the copy stream (240 MB = 4.5 ms on K5) and hold
// back every kernel of the call -- measured: flash_cost_grad 21.9 ms alone, 26.7 ms with the next frame uploading.
__global__ void pull_doubles_kernel(double* __restrict__ dst, const double* __restrict__ src, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
int pull_pinned(flash_ctx* ctx, double* dst, const double* src_pinned, size_t n) {
  if (n == 0) return 0;
  pull_doubles_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 64), 256, 0, ctx->stream>>>(dst, src_pinned, n);
  CU(cudaGetLastError());
  return 0;
}

SmemLayout cost_layout(const flash_ctx* ctx, bool grad) {
  const ModelC& M = ctx->hm.M;
  return smem_layout(M.tab_len, M.acc_len, M.tabf_len, grad && M.n_skins > 0, CWARPS);
}

template <bool G, bool A, bool D, bool DET = false>
int launch_cost_t(flash_ctx* ctx, const CostParams& p, int C, int tiles, size_t smem) {
  auto k = cost_kernel<G, A, D, DET>;
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)C, (unsigned)tiles);
  k<<<grid, CBLOCK, smem, ctx->stream>>>(p);
  CU(cudaGetLastError());
  ctx->launches++; ctx->last.launches++;
  return 0;
}

// points per block: enough blocks for >= ~4 waves, at most 4096 points, at least one batch
int choose_tile(const flash_ctx* ctx, long long n, int C, bool lpt = false) {
  // small problems (tracking: one config, one depth frame): fewer full-size blocks than SMs would leave SMs idle
  // while the others queue 12 warps on their fp64 pipe; spread the points over one block per SM instead, in whole
  // warp groups of 32 (a block's last batch is then partial)
  if ((n * C + CBLOCK - 1) / CBLOCK < 2ll * ctx->sm_count && n > 0) {
    const long long tiles = std::max<long long>(1, ctx->sm_count / std::max(C, 1));
    long long ppb = (n + tiles - 1) / tiles;
    ppb = std::max<long long>(32, ((ppb + 31) / 32) * 32);
    return (int)ppb;
  }
  // >= 20 waves of blocks (2 per SM) where there is that much work (swept 4 .. 64 on K4 and on K5 at the per-rank size of 8 GPUs): the blocks' durations differ several-fold (tiles near the
  // bodies), so few waves end in a long tail -- K4's 8 x 16 batch ran as 1,280 blocks = 4.3 waves with 32,768-point tiles
  // With the heaviest tiles launched first (tile_order_kernel) the tail is short anyway: half as many, larger blocks (K5 shard of
  // 1.25 M points: 20 waves 2.520 ms, 10 waves 2.496; full cloud, 65,536-point tiles: 19.52 against 19.66 ms)
  static const long long waves_env = [] { const char* e = getenv("FLASHSDF_WAVES"); long long v = e ? atoll(e) : 0; return v > 0 ? v : 0ll; }();
  const long long waves = waves_env ? waves_env : (lpt ? 10ll : 20ll);
  const long long target_blocks = (long long)ctx->sm_count * 2 * waves;
  long long ppb = (n * C + target_blocks - 1) / target_blocks;
  ppb = ((ppb + CBLOCK - 1) / CBLOCK) * CBLOCK;
  if (ppb < CBLOCK) ppb = CBLOCK;
  static const long long max_tile = [] { const char* e = getenv("FLASHSDF_TILE"); long long v = e ? (atoll(e) / 32) * 32 : 0; return v >= CBLOCK ? v : (32768ll / CBLOCK) * CBLOCK; }();   // K5: 4032 -> 55.5 ms, 8064 -> 54.0, 16128 -> 53.5, 32256 -> 53.5; with the group pre-pass 16128 -> 45.4, 24576 -> 44.6, 32256 -> 44.5
  if (ppb > (lpt ? 2 * max_tile : max_tile)) ppb = lpt ? 2 * max_tile : max_tile;
  while ((n + ppb - 1) / ppb > 65535) ppb *= 2;
  return (int)ppb;
}

// cost_kernel's tiles are launched heaviest first where group_kernel runs anyway (its records give the estimates) and there
// are enough (group, configuration) pairs for a tail to matter.  A batch of frames gets one order per frame (one order for
// all frames mixes places in a block row: K4 measured 3 % slower with that, 2 % faster with its own order per frame).
static bool lpt_applies(const flash_ctx* ctx, int C, int cfg_per_frame) {
  static const int mode = [] { const char* e = getenv("FLASHSDF_TILE_ORDER"); return e ? atoi(e) : 2; }();   // 0: off, 1: single frames only, 2: batches of frames too
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  const long long n_tile = cfg_per_frame > 0 ? P.max_cnt : P.n;
  return mode > 0 && (cfg_per_frame == 0 || mode > 1) && P.grp_on && ctx->group_cull && ctx->hm.M.n_convex >= 2 && ((n_tile + 31) / 32) * C >= 262144;
}
// group pre-pass (group_kernel) for C configurations of the current point set -> ctx->d_rec; *stride = records per config
int run_groups(flash_ctx* ctx, int C, int cfg_per_frame, long long* stride, int pts_per_block = 0, bool* ordered = nullptr) {
  const ModelC& M = ctx->hm.M;
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  *stride = 0;
  const long long n_tile = cfg_per_frame > 0 ? P.max_cnt : P.n;
  if (!(P.grp_on && ctx->group_cull) || M.n_convex < 2 || n_tile == 0) return 0;
  const long long G = (n_tile + 31) / 32;
  int rc = ensure(ctx, ctx->d_rec, ctx->cap_rec, (size_t)C * (size_t)G); if (rc) return rc;
  GroupParams g;
  g.hot = ctx->d_hot; g.geo_f4 = ctx->d_geo_f4; g.geo_u32 = ctx->d_geo_u32;
  g.tab = ctx->d_tab; g.grp = P.d_grp; g.frames = cfg_per_frame > 0 ? P.d_frames : nullptr; g.cfg_per_frame = cfg_per_frame > 0 ? cfg_per_frame : 1;
  g.n_groups = G; g.rec = ctx->d_rec; g.rec_stride = G;
  // groups per block: >= 16 waves of blocks when there is that much work (one thread per group: a block's cost varies with
  // the groups' candidate counts, so few big blocks end in a long tail -- 1,280 blocks of 2,048 groups cost K5 at 1.25 M
  // points 0.35 ms against 0.25 ms pro rata), never less than one group per thread
  long long gpb = (G * C + (long long)ctx->sm_count * 8 * 16 - 1) / ((long long)ctx->sm_count * 8 * 16);
  gpb = std::min<long long>(std::max<long long>(((gpb + GBLOCK - 1) / GBLOCK) * GBLOCK, GBLOCK), 8 * GBLOCK);
  g.groups_per_block = (int)gpb;
  // exact centre distances pay when there are enough (group, configuration) pairs to fill the GPU with one thread each:
  // K5 (2e7 pairs) 28.6 -> 22.0 ms; on a single depth frame (K2: 9,600 pairs) the longer serial threads cost more than
  // the tighter candidate sets save (0.069 -> 0.088 ms)
  g.exact = (ctx->exact_groups && !ctx->use_grid && G * C >= 262144) ? 1 : 0; g.want_cull = 1;
  // tile order of cost_kernel (heaviest first) where there are enough blocks for a tail to matter and few enough tiles to sort
  const long long tiles = pts_per_block > 0 ? (n_tile + pts_per_block - 1) / pts_per_block : 0;
  g.tile_w = nullptr; g.groups_per_tile = 1; g.tiles = 0;
  if (lpt_applies(ctx, C, cfg_per_frame) && tiles >= 8 && tiles <= 1024) {
    const size_t slots = (size_t)MAX_FRAMES * 1024;
    if (!ctx->d_tile_w) { CU(cudaMalloc((void**)&ctx->d_tile_w, slots * sizeof(unsigned))); CU(cudaMemsetAsync(ctx->d_tile_w, 0, slots * sizeof(unsigned), ctx->stream)); }
    rc = ensure(ctx, ctx->d_order, ctx->cap_order, slots); if (rc) return rc;
    g.tile_w = ctx->d_tile_w; g.groups_per_tile = pts_per_block / 32; g.tiles = (int)tiles;
  }
  const size_t smem = group_smem(M.tabf_len);
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  group_kernel<<<dim3((unsigned)C, (unsigned)((G + gpb - 1) / gpb)), GBLOCK, smem, ctx->stream>>>(g);
  CU(cudaGetLastError());
  ctx->launches++; ctx->last.launches++;
  if (g.tile_w) {
    int nt = 32; while (nt < tiles) nt <<= 1;
    tile_order_kernel<<<cfg_per_frame > 0 ? P.ft.n_frames : 1, nt, 0, ctx->stream>>>(ctx->d_tile_w, (int)tiles, ctx->d_order);
    CU(cudaGetLastError());
    ctx->launches++; ctx->last.launches++;
    if (ordered) *ordered = true;
  }
  *stride = G;
  return 0;
}

// C = total configurations of the launch; cfg_per_frame > 0: configuration c evaluates frame c / cfg_per_frame only
int run_cost(flash_ctx* ctx, int C, bool grad, int32_t* d_argmin, int cfg_per_frame = 0) {
  const ModelC& M = ctx->hm.M;
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  long long rec_stride = 0;
  bool ordered = false;
  const int ppb = choose_tile(ctx, cfg_per_frame > 0 ? P.max_cnt : P.n, C, lpt_applies(ctx, C, cfg_per_frame));
  { int rc = run_groups(ctx, C, cfg_per_frame, &rec_stride, ppb, &ordered); if (rc) return rc; }
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[5], ctx->stream));
  CostParams p;
  p.rec = rec_stride > 0 ? ctx->d_rec : nullptr; p.rec_stride = rec_stride;
  p.hot = ctx->d_hot; p.geo_f4 = ctx->d_geo_f4; p.geo_u32 = ctx->d_geo_u32; p.grid_pool = ctx->d_grid_pool;
  p.pts = P.d_pts; p.n_pts = P.n; p.tab = ctx->d_tab; p.acc = ctx->d_acc; p.argmin = d_argmin; p.dist = nullptr; p.perm = P.perm_on ? P.d_perm : nullptr;
  p.grp = P.grp_on && ctx->group_cull ? P.d_grp : nullptr;
  p.frames = cfg_per_frame > 0 ? P.d_frames : nullptr; p.cfg_per_frame = cfg_per_frame > 0 ? cfg_per_frame : 1;
  const long long n_tile = cfg_per_frame > 0 ? P.max_cnt : P.n;
  p.pts_per_block = ppb;
  p.use_grid = ctx->use_grid ? 1 : 0;
  if (P.n == 0 || n_tile == 0) return 0;
  const int tiles = (int)((n_tile + p.pts_per_block - 1) / p.pts_per_block);
  const size_t smem = cost_layout(ctx, grad).total;
  if (smem + COST_STATIC_SMEM > (size_t)ctx->smem_optin) return fail(ctx, FLASH_ERR_TOO_LARGE, "model tables do not fit in shared memory (" + std::to_string(smem) + " bytes)");
  p.part = nullptr; p.tile_order = nullptr;
  if (ordered) p.tile_order = ctx->d_order;
  if (ctx->deterministic) {
    int rc = ensure(ctx, ctx->d_part, ctx->cap_part, (size_t)C * (size_t)tiles * (size_t)M.acc_len); if (rc) return rc;
    p.part = ctx->d_part;
  }
  int rc;
  if (ctx->deterministic) {
    if (grad) rc = d_argmin ? launch_cost_t<true, true, false, true>(ctx, p, C, tiles, smem) : launch_cost_t<true, false, false, true>(ctx, p, C, tiles, smem);
    else rc = d_argmin ? launch_cost_t<false, true, false, true>(ctx, p, C, tiles, smem) : launch_cost_t<false, false, false, true>(ctx, p, C, tiles, smem);
  } else if (grad) rc = d_argmin ? launch_cost_t<true, true, false>(ctx, p, C, tiles, smem) : launch_cost_t<true, false, false>(ctx, p, C, tiles, smem);
  else rc = d_argmin ? launch_cost_t<false, true, false>(ctx, p, C, tiles, smem) : launch_cost_t<false, false, false>(ctx, p, C, tiles, smem);
  if (rc || !ctx->deterministic) return rc;
  reduce_partials_kernel<<<C, 128, 0, ctx->stream>>>(ctx->d_part, ctx->d_acc, M.acc_len, tiles, p.frames, p.cfg_per_frame, P.n, p.pts_per_block, grad ? M.acc_len : 1);
  CU(cudaGetLastError());
  ctx->launches++; ctx->last.launches++;
  return 0;
}

int run_prepare(flash_ctx* ctx, int C, bool zero_acc = false) {
  const size_t smem = prepare_smem_doubles(ctx->hm.M) * sizeof(double);
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(prepare_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  prepare_kernel<<<C, PREP_T, smem, ctx->stream>>>(ctx->d_model, ctx->d_X, ctx->d_tab, ctx->d_aux, zero_acc ? ctx->d_acc : nullptr, ctx->d_status);
  CU(cudaGetLastError());
  ctx->launches++; ctx->last.launches++;
  return 0;
}

int run_epilogue(flash_ctx* ctx, int C, double w_def, bool grad) {
  const size_t smem = epilogue_smem_doubles(ctx->hm.M) * sizeof(double);
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(epilogue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  epilogue_kernel<<<C, 32, smem, ctx->stream>>>(ctx->d_model, ctx->d_X, ctx->d_tab, ctx->d_aux, ctx->d_acc, C, w_def, ctx->d_cost, grad ? ctx->d_grad : nullptr);
  CU(cudaGetLastError());
  ctx->launches++; ctx->last.launches++;
  return 0;
}

int run_allreduce(flash_ctx* ctx, int C) {
  if (!ctx->comm) return 0;
  int rc = g_nccl.AllReduce(ctx->d_acc, ctx->d_acc, (size_t)C * ctx->hm.M.acc_len, NCCL_FLOAT64, NCCL_SUM, ctx->comm, ctx->stream);
  if (rc != 0) return fail(ctx, FLASH_ERR_NCCL, std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"));
  return 0;
}

int collect_timing(flash_ctx* ctx) {
  if (!ctx->profiling) return 0;
  CU(cudaEventSynchronize(ctx->ev[4]));
  CU(cudaEventElapsedTime(&ctx->last.ms_prepare, ctx->ev[0], ctx->ev[1]));
  CU(cudaEventElapsedTime(&ctx->last.ms_main, ctx->ev[1], ctx->ev[2]));
  CU(cudaEventElapsedTime(&ctx->last.ms_group, ctx->ev[1], ctx->ev[5]));
  CU(cudaEventElapsedTime(&ctx->last.ms_allreduce, ctx->ev[2], ctx->ev[3]));
  CU(cudaEventElapsedTime(&ctx->last.ms_epilogue, ctx->ev[3], ctx->ev[4]));
  CU(cudaEventElapsedTime(&ctx->last.ms_total, ctx->ev[0], ctx->ev[4]));
  return 0;
}

int check_status(flash_ctx* ctx) {
  int st = 0;
  CU(cudaMemcpyAsync(&st, ctx->d_status, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (st != 0) {
    CU(cudaMemsetAsync(ctx->d_status, 0, sizeof(int), ctx->stream));
    if (st == 2) return fail(ctx, FLASH_ERR_CUDA, "track_kernel: the grid barrier timed out (blocks not co-resident); set FLASHSDF_NO_TRACK_KERNEL=1");
    return fail(ctx, FLASH_ERR_INVALID, "singular RBF system (coincident skin control points)");
  }
  return 0;
}

// grid mode: allocate + bake body-frame SDF grids for every convex surface
int bake_grids(flash_ctx* ctx) {
  ModelC& M = ctx->hm.M;
  const double h = ctx->hm.opt.grid_spacing, margin = ctx->hm.opt.grid_margin;
  size_t total = 0;
  for (int j = 0; j < M.n_convex; ++j) {
    ConvexC& C = M.cvx[j];
    long long cells = 1;
    for (int k = 0; k < 3; ++k) {
      C.g_org[k] = (float)(C.lo[k] - margin);
      C.g_dim[k] = (int)std::ceil((C.hi[k] - C.lo[k] + 2 * margin) / h) + 1;
      cells *= C.g_dim[k];
    }
    if (cells > (1ll << 31)) return fail(ctx, FLASH_ERR_TOO_LARGE, "SDF grid too large; increase grid_spacing");
    C.g_inv_h = (float)(1.0 / h);
    C.g_off = (long long)total;
    C.g_on = 1;
    total += (size_t)cells;
  }
  if (total * sizeof(float) > (size_t)64 << 30) return fail(ctx, FLASH_ERR_TOO_LARGE, "SDF grids exceed 64 GiB; increase grid_spacing");
  CU(cudaMalloc((void**)&ctx->d_grid_pool, std::max<size_t>(total, 1) * sizeof(float)));
  ctx->grid_pool_floats = total;
  make_hot(M, ctx->hot);
  CU(cudaMemcpy(ctx->d_model, &M, sizeof(ModelC), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(ctx->d_hot, &ctx->hot, sizeof(HotModel), cudaMemcpyHostToDevice));
  unsigned int* d_err = nullptr;
  CU(cudaMalloc((void**)&d_err, sizeof(unsigned int) * std::max(M.n_convex, 1)));
  CU(cudaMemset(d_err, 0, sizeof(unsigned int) * std::max(M.n_convex, 1)));
  for (int j = 0; j < M.n_convex; ++j) {
    const ConvexC& C = M.cvx[j];
    long long n = (long long)C.g_dim[0] * C.g_dim[1] * C.g_dim[2];
    bake_grid_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_hot, ctx->d_geo_f4, ctx->d_geo_u32, j, ctx->d_grid_pool);
    CU(cudaGetLastError());
    long long nc = (long long)(C.g_dim[0] - 1) * (C.g_dim[1] - 1) * (C.g_dim[2] - 1);
    grid_error_kernel<<<(unsigned)((nc + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_hot, ctx->d_geo_f4, ctx->d_geo_u32, j, ctx->d_grid_pool, d_err + j);
    CU(cudaGetLastError());
    ctx->launches += 2;
  }
  std::vector<unsigned int> bits(std::max(M.n_convex, 1));
  CU(cudaMemcpyAsync(bits.data(), d_err, sizeof(unsigned int) * bits.size(), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  CU(cudaFree(d_err));
  ctx->grid_err.assign(M.n_convex, 0.0);
  for (int j = 0; j < M.n_convex; ++j) { float f; std::memcpy(&f, &bits[j], 4); ctx->grid_err[j] = f; }
  ctx->use_grid = true;
  return 0;
}

}  // namespace

// ---- point sets: upload, Hilbert sort, group balls (flash_ingest.cuh), all on the copy stream -------------------
static int set_alloc(flash_ctx* ctx, flash_ctx::PointSet& P, long long span) {
  if (P.external) { P.d_pts = nullptr; P.cap_pts = 0; P.external = false; }
  if (span > P.cap_pts) {
    if (P.d_pts) CU(cudaFree(P.d_pts));
    P.d_pts = nullptr; P.cap_pts = 0;
    const long long want = span + span / 16 + 32;
    CU(cudaMalloc((void**)&P.d_pts, (size_t)want * sizeof(float4)));
    P.cap_pts = want;
  }
  if (!P.d_frames) CU(cudaMalloc((void**)&P.d_frames, (size_t)(2 * MAX_FRAMES + 2) * sizeof(long long)));
  if (!P.ready) { CU(cudaEventCreateWithFlags(&P.ready, cudaEventDisableTiming)); CU(cudaEventCreateWithFlags(&P.retired, cudaEventDisableTiming)); }
  return 0;
}

static bool host_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// F frames ptrs[f][counts[f]][3] (host) -> set P.  Returns after ENQUEUEING the work on the copy stream; pageable
// caller memory has been staged into the ctx's pinned ring by then, pinned caller memory is read by DMA until P.ready.
// segments = true: the F inputs are pieces of ONE frame (a multi-device parent hands every device several chunks of the
// caller's cloud), concatenated in the staging buffer.
template <class T>
static int ingest(flash_ctx* ctx, flash_ctx::PointSet& P, const T* const* ptrs, const int64_t* counts, int F, bool segments = false) {
  CU(cudaSetDevice(ctx->device));
  FrameTable& ft = P.ft;
  long long n = 0, span = 0, max_cnt = 0;
  std::vector<long long> in_off((size_t)F + 1, 0);   // offset of every input in the caller-order staging buffer
  for (int f = 0; f < F; ++f) { in_off[f] = n; n += counts[f]; }
  in_off[F] = n;
  const int NF = segments ? 1 : F;
  ft.n_frames = NF;
  if (segments) {
    ft.src_off[0] = 0; ft.off[0] = 0; ft.cnt[0] = n; span = (n + 31) & ~31ll; max_cnt = n;
  } else {
    for (int f = 0; f < F; ++f) {
      ft.src_off[f] = in_off[f]; ft.off[f] = span; ft.cnt[f] = counts[f];
      span += (counts[f] + 31) & ~31ll; max_cnt = std::max<long long>(max_cnt, counts[f]);
    }
  }
  ft.src_off[NF] = n; ft.off[NF] = span;
  int rc = set_alloc(ctx, P, std::max<long long>(span, 32)); if (rc) return rc;
  cudaStream_t cs = ctx->copy_stream;
  if (P.retired_valid) { CU(cudaStreamWaitEvent(cs, P.retired, 0)); P.retired_valid = false; }
  P.perm_on = false; P.grp_on = false; P.n = n; P.max_cnt = max_cnt; P.set = true;
  {
    long long h[2 * MAX_FRAMES + 2];
    for (int f = 0; f <= NF; ++f) h[f] = ft.off[f];
    for (int f = 0; f < NF; ++f) h[MAX_FRAMES + 1 + f] = ft.cnt[f];
    // small and rare (once per frame batch): pageable source, synchronous staging by the driver
    CU(cudaMemcpyAsync(P.d_frames, h, sizeof h, cudaMemcpyHostToDevice, cs));
  }
  if (n == 0) { CU(cudaEventRecord(P.ready, cs)); return FLASH_OK; }
  // ---- upload into the caller-order staging buffer
  const size_t bytes = (size_t)n * 3 * sizeof(T);
  if (bytes > ctx->cap_stage) {
    CU(cudaStreamSynchronize(cs));
    if (ctx->d_stage) CU(cudaFree(ctx->d_stage));
    ctx->d_stage = nullptr; ctx->cap_stage = 0;
    CU(cudaMalloc(&ctx->d_stage, bytes + bytes / 16));
    ctx->cap_stage = bytes + bytes / 16;
  }
  bool all_pinned = true;
  for (int f = 0; f < F; ++f) if (counts[f] > 0 && !host_pinned(ptrs[f])) all_pinned = false;
  if (all_pinned) {
    for (int f = 0; f < F; ++f)
      if (counts[f] > 0) CU(cudaMemcpyAsync((char*)ctx->d_stage + (size_t)in_off[f] * 3 * sizeof(T), ptrs[f], (size_t)counts[f] * 3 * sizeof(T), cudaMemcpyHostToDevice, cs));
  } else {
    const int slot = ctx->ring_next; ctx->ring_next ^= 1;
    if (!ctx->ring_done[slot]) CU(cudaEventCreateWithFlags(&ctx->ring_done[slot], cudaEventDisableTiming));
    else CU(cudaEventSynchronize(ctx->ring_done[slot]));          // the DMA that last read this slot has finished
    if (bytes > ctx->cap_ring[slot]) {
      if (ctx->h_ring[slot]) CU(cudaFreeHost(ctx->h_ring[slot]));
      ctx->h_ring[slot] = nullptr; ctx->cap_ring[slot] = 0;
      CU(cudaHostAlloc(&ctx->h_ring[slot], bytes + bytes / 16, cudaHostAllocDefault));
      ctx->cap_ring[slot] = bytes + bytes / 16;
    }
    for (int f = 0; f < F; ++f)
      if (counts[f] > 0) std::memcpy((char*)ctx->h_ring[slot] + (size_t)in_off[f] * 3 * sizeof(T), ptrs[f], (size_t)counts[f] * 3 * sizeof(T));
    CU(cudaMemcpyAsync(ctx->d_stage, ctx->h_ring[slot], bytes, cudaMemcpyHostToDevice, cs));
    CU(cudaEventRecord(ctx->ring_done[slot], cs));
  }
  const T* src = (const T*)ctx->d_stage;
  const long long span_groups = span >> 5;
  rc = ensure(ctx, P.d_grp, P.cap_grp, (size_t)span_groups); if (rc) return rc;
  const unsigned gblocks = (unsigned)((span_groups * 32 + 255) / 256);
  if (ctx->sort_points && n >= 4096 && n < (1ll << 31)) {
    rc = ensure(ctx, P.d_perm, P.cap_perm, (size_t)span); if (rc) return rc;
    if ((size_t)n > ctx->cap_sort) {
      CU(cudaStreamSynchronize(cs));
      for (int k = 0; k < 2; ++k) { if (ctx->d_keys[k]) CU(cudaFree(ctx->d_keys[k])); if (ctx->d_idx[k]) CU(cudaFree(ctx->d_idx[k])); ctx->d_keys[k] = nullptr; ctx->d_idx[k] = nullptr; }
      const size_t want = (size_t)n + (size_t)n / 16;
      for (int k = 0; k < 2; ++k) { CU(cudaMalloc((void**)&ctx->d_keys[k], want * sizeof(unsigned))); CU(cudaMalloc((void**)&ctx->d_idx[k], want * sizeof(int))); }
      ctx->cap_sort = want;
    }
    if (!ctx->d_bounds) CU(cudaMalloc((void**)&ctx->d_bounds, (size_t)(6 * MAX_FRAMES + 1) * sizeof(unsigned)));
    unsigned* scan_done = ctx->d_bounds + 6 * MAX_FRAMES;
    // key layout: frame bits | 3 * hb Hilbert bits; the top `kb` bits are sorted (enough to order clouds of this size)
    int fb = 0; while ((1 << fb) < NF) ++fb;
    const int hb = std::min(10, (32 - fb) / 3);
    const int total_bits = fb + 3 * hb;
    int lg = 0; while ((1ll << lg) < n) ++lg;
    int kb = std::min(total_bits, std::max(16, lg + 5 + fb));
    const int passes = (kb + 7) / 8;
    kb = std::min(total_bits, passes * 8);
    const int drop = total_bits - kb;
    // chunk of keys per warp: a multiple of 32, about 4096 warps in flight for large clouds
    long long E = ((n + 4095) / 4096 + 31) & ~31ll;
    E = std::min<long long>(std::max<long long>(E, 512), 1 << 20);
    const int nW = (int)((n + E - 1) / E);
    const long long tlen = 256ll * nW;
    const int sblocks = (int)((tlen + 1023) / 1024);
    if ((size_t)tlen > ctx->cap_table) {
      CU(cudaStreamSynchronize(cs));
      if (ctx->d_table) CU(cudaFree(ctx->d_table));
      if (ctx->d_sums) CU(cudaFree(ctx->d_sums));
      ctx->d_table = nullptr; ctx->d_sums = nullptr; ctx->cap_table = 0;
      CU(cudaMalloc((void**)&ctx->d_table, (size_t)tlen * 2 * sizeof(int)));
      CU(cudaMalloc((void**)&ctx->d_sums, ((size_t)tlen * 2 / 1024 + 2) * sizeof(int)));
      ctx->cap_table = (size_t)tlen * 2;
    }
    bounds_init_kernel<<<(6 * NF + 255) / 256, 256, 0, cs>>>(ctx->d_bounds, NF, scan_done);
    const unsigned bb = (unsigned)std::min<long long>((max_cnt + 255) / 256, std::max(1, ctx->sm_count * 16 / NF));
    frames_bounds_kernel<T><<<dim3(std::max(bb, 1u), NF), 256, 0, cs>>>(src, ft, ctx->d_bounds);
    points_keys_kernel<T><<<dim3((unsigned)((max_cnt + 255) / 256), NF), 256, 0, cs>>>(src, ft, ctx->d_bounds, hb, drop, ctx->d_keys[0], ctx->d_idx[0]);
    const unsigned rblocks = (unsigned)((nW + RS_WARPS - 1) / RS_WARPS);
    int in = 0;
    for (int ps = 0; ps < passes; ++ps) {
      radix_hist_kernel<<<rblocks, RS_BLOCK, 0, cs>>>(ctx->d_keys[in], n, 8 * ps, (int)E, nW, ctx->d_table);
      radix_scan_kernel<<<sblocks, 1024, 0, cs>>>(ctx->d_table, tlen, ctx->d_sums, scan_done);
      radix_scatter_kernel<<<rblocks, RS_BLOCK, 0, cs>>>(ctx->d_keys[in], ctx->d_idx[in], n, 8 * ps, (int)E, nW, ctx->d_table, ctx->d_sums, ctx->d_keys[in ^ 1], ctx->d_idx[in ^ 1]);
      in ^= 1;
    }
    points_gather_kernel<T><<<gblocks, 256, 0, cs>>>(src, ctx->d_idx[in], ft, P.d_pts, P.d_perm, P.d_grp);
    CU(cudaGetLastError());
    ctx->launches += 4 + 3 * passes;
    P.perm_on = true;
  } else {
    points_convert_kernel<T><<<gblocks, 256, 0, cs>>>(src, ft, P.d_pts, P.d_grp);
    CU(cudaGetLastError());
    ctx->launches++;
  }
  P.grp_on = true;
  CU(cudaEventRecord(P.ready, cs));
  return FLASH_OK;
}

// back set -> front: kernels issued from now on read the new frame(s); the old front may be overwritten by the next
// upload once everything already issued on the compute stream has run
static int swap_sets(flash_ctx* ctx) {
  flash_ctx::PointSet& old = ctx->ps[ctx->cur];
  if (old.ready) { CU(cudaEventRecord(old.retired, ctx->stream)); old.retired_valid = true; }
  ctx->cur ^= 1;
  ctx->back_pending = false;
  ctx->n_global = -1.0;
  CU(cudaStreamWaitEvent(ctx->stream, ctx->ps[ctx->cur].ready, 0));
  return FLASH_OK;
}

template <class T>
static int points_set_sync(flash_ctx* ctx, const T* xyz, int64_t n) {
  const T* ptrs[1] = {xyz};
  const int64_t counts[1] = {n};
  int rc = ingest<T>(ctx, ctx->ps[ctx->cur ^ 1], ptrs, counts, 1); if (rc) return rc;
  rc = swap_sets(ctx); if (rc) return rc;
  CU(cudaEventSynchronize(ctx->ps[ctx->cur].ready));   // the caller may reuse xyz as soon as we return
  return FLASH_OK;
}

// multi-device forms (defined further down, next to the staged cost_grad they are built from)
template <class T> static int multi_points_set(flash_ctx* ctx, const T* xyz, int64_t n, bool async);
static int multi_swap(flash_ctx* ctx);
static int multi_frames_set(flash_ctx* ctx, const double* const* xyz, const int64_t* counts, int32_t F, bool async);
static int multi_create(const flash_model_desc* desc, const flash_options* opt, flash_ctx** out);

// ================================================================================================
extern "C" {

void flash_options_default(flash_options* opt) { if (opt) options_default(opt); }
int flash_abi_version(void) { return FLASHSDF_ABI_VERSION; }
const char* flash_last_error(const flash_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int flash_model_create(const flash_model_desc* desc, const flash_options* opt, flash_ctx** out) {
  if (!out) return fail(nullptr, FLASH_ERR_INVALID, "null out pointer");
  *out = nullptr;
  if (opt && opt->abi_version == FLASHSDF_ABI_VERSION && opt->n_devices > 1) return multi_create(desc, opt, out);
  flash_ctx* ctx = new flash_ctx;
  std::string err;
  if (!build_model(desc, opt, ctx->hm, err)) { delete ctx; return fail(nullptr, FLASH_ERR_INVALID, err); }
  make_hot(ctx->hm.M, ctx->hot);
  if (const char* e = getenv("FLASHSDF_NO_SORT")) ctx->sort_points = !(e[0] == '1');
  if (const char* e = getenv("FLASHSDF_NO_GROUP_CULL")) ctx->group_cull = !(e[0] == '1');
  if (const char* e = getenv("FLASHSDF_NO_EXACT_GROUPS")) ctx->exact_groups = !(e[0] == '1');
  ctx->deterministic = ctx->hm.opt.deterministic != 0;
  if (const char* e = getenv("FLASHSDF_DETERMINISTIC")) ctx->deterministic = (e[0] == '1');
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    delete ctx;
    return fail(nullptr, FLASH_ERR_NO_DEVICE, std::string("no CUDA device (libflashsdf has no CPU path): ") + cudaGetErrorString(e));
  }
  int dev = ctx->hm.opt.device;
  if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) dev = 0; }
  if (dev >= ndev) { delete ctx; return fail(nullptr, FLASH_ERR_INVALID, "device ordinal out of range"); }
  ctx->device = dev;
  auto bail = [&](int rc) { g_create_err = ctx->err; flash_destroy(ctx); return rc; };
#define CUC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_); return bail(FLASH_ERR_CUDA); } } while (0)
  CUC(cudaSetDevice(dev));
  cudaDeviceProp prop;
  CUC(cudaGetDeviceProperties(&prop, dev));
  ctx->sm_count = prop.multiProcessorCount; ctx->l2_bytes = prop.l2CacheSize; ctx->smem_optin = (int)prop.sharedMemPerBlockOptin;
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev); ctx->clock_khz = khz;
  CUC(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  CUC(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  for (auto& ev : ctx->ev) CUC(cudaEventCreate(&ev));
  const ModelC& M = ctx->hm.M;
  CUC(cudaMalloc((void**)&ctx->d_model, sizeof(ModelC)));
  CUC(cudaMalloc((void**)&ctx->d_hot, sizeof(HotModel)));
  CUC(cudaMalloc((void**)&ctx->d_geo_f4, std::max<size_t>(M.geo_f4, 1) * sizeof(F4)));
  CUC(cudaMalloc((void**)&ctx->d_geo_u32, std::max<size_t>(M.geo_u32, 1) * sizeof(uint32_t)));
  CUC(cudaMalloc((void**)&ctx->d_status, sizeof(int)));
  CUC(cudaMemset(ctx->d_status, 0, sizeof(int)));
  CUC(cudaMemcpy(ctx->d_model, &M, sizeof(ModelC), cudaMemcpyHostToDevice));
  CUC(cudaMemcpy(ctx->d_hot, &ctx->hot, sizeof(HotModel), cudaMemcpyHostToDevice));
  if (M.geo_f4) CUC(cudaMemcpy(ctx->d_geo_f4, ctx->hm.geo_f4.data(), M.geo_f4 * sizeof(F4), cudaMemcpyHostToDevice));
  if (M.geo_u32) CUC(cudaMemcpy(ctx->d_geo_u32, ctx->hm.geo_u32.data(), M.geo_u32 * sizeof(uint32_t), cudaMemcpyHostToDevice));
  const size_t smem = cost_layout(ctx, true).total;
  if (smem + COST_STATIC_SMEM > (size_t)ctx->smem_optin) { ctx->err = "model geometry does not fit in shared memory (" + std::to_string(smem) + " bytes)"; return bail(FLASH_ERR_TOO_LARGE); }
  if (ctx->hm.opt.sdf_mode == FLASH_SDF_GRID) { int rc = bake_grids(ctx); if (rc) return bail(rc); }
#undef CUC
  *out = ctx;
  return FLASH_OK;
}

void flash_destroy(flash_ctx* ctx) {
  if (!ctx) return;
  if (!ctx->subs.empty()) {
    for (flash_ctx* sub : ctx->subs) flash_destroy(sub);
    delete ctx;
    return;
  }
  if (ctx->d_peer_acc) { cudaSetDevice(ctx->device); cudaFree(ctx->d_peer_acc); }
  if (ctx->ev_cost) cudaEventDestroy(ctx->ev_cost);
  if (ctx->ev_red) cudaEventDestroy(ctx->ev_red);
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(ctx->comm);
  if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
  void* frees[] = {ctx->d_model, ctx->d_hot, ctx->d_geo_f4, ctx->d_geo_u32, ctx->d_grid_pool, ctx->d_stage,
                   ctx->d_X, ctx->d_tab, ctx->d_aux, ctx->d_acc, ctx->d_cost, ctx->d_grad, ctx->d_div, ctx->d_status, ctx->d_done, ctx->d_iters,
                   ctx->d_argmin, ctx->d_rec, ctx->d_part, ctx->d_bar, ctx->d_order, ctx->d_tile_w, ctx->d_qd, ctx->d_qi, ctx->d_keys[0], ctx->d_keys[1], ctx->d_idx[0], ctx->d_idx[1], ctx->d_table, ctx->d_sums, ctx->d_bounds};
  for (void* p : frees) if (p) cudaFree(p);
  for (auto& P : ctx->ps) {
    if (P.d_pts && !P.external) cudaFree(P.d_pts);
    if (P.d_perm) cudaFree(P.d_perm);
    if (P.d_grp) cudaFree(P.d_grp);
    if (P.d_frames) cudaFree(P.d_frames);
    if (P.ready) cudaEventDestroy(P.ready);
    if (P.retired) cudaEventDestroy(P.retired);
  }
  for (int k = 0; k < 2; ++k) { if (ctx->h_ring[k]) cudaFreeHost(ctx->h_ring[k]); if (ctx->ring_done[k]) cudaEventDestroy(ctx->ring_done[k]); }
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  if (ctx->h_pin) cudaFreeHost(ctx->h_pin);
  for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

// ---- URDF / mesh / point-cloud files (flash_urdf.h; no CUDA involved in parsing) ----
struct flash_urdf { flashurdf::Urdf u; };

int flash_urdf_load(const char* urdf_path, const char* const* package_dirs, int32_t n_package_dirs, flash_urdf** out) {
  if (!out) return fail(nullptr, FLASH_ERR_INVALID, "flash_urdf_load: null out pointer");
  *out = nullptr;
  if (!urdf_path || n_package_dirs < 0 || (n_package_dirs > 0 && !package_dirs)) return fail(nullptr, FLASH_ERR_INVALID, "flash_urdf_load: bad arguments");
  std::vector<std::string> dirs;
  for (int i = 0; i < n_package_dirs; ++i) if (package_dirs[i]) dirs.push_back(package_dirs[i]);
  flash_urdf* h = new flash_urdf;
  std::string err;
  if (!flashurdf::load(urdf_path, dirs, h->u, err)) { delete h; return fail(nullptr, FLASH_ERR_IO, err); }
  *out = h;
  return FLASH_OK;
}
const flash_model_desc* flash_urdf_desc(const flash_urdf* u) { return u ? &u->u.desc : nullptr; }
const char* flash_urdf_body_name(const flash_urdf* u, int32_t body) { return (u && body >= 0 && body < (int32_t)u->u.body_name.size()) ? u->u.body_name[body].c_str() : nullptr; }
const char* flash_urdf_surface_mesh(const flash_urdf* u, int32_t surface) { return (u && surface >= 0 && surface < (int32_t)u->u.surface_mesh.size()) ? u->u.surface_mesh[surface].c_str() : nullptr; }
int flash_urdf_set_joint_type(flash_urdf* u, int32_t body, int32_t joint_type) {
  if (!u || body < 0 || body >= (int32_t)u->u.joint_type.size() || joint_type < 0 || joint_type > 2) return fail(nullptr, FLASH_ERR_INVALID, "flash_urdf_set_joint_type: bad arguments");
  u->u.joint_type[body] = joint_type;
  return FLASH_OK;
}
void flash_urdf_free(flash_urdf* u) { delete u; }
int flash_model_load_urdf(const char* urdf_path, const char* const* package_dirs, int32_t n_package_dirs, const flash_options* opt, flash_ctx** out) {
  flash_urdf* u = nullptr;
  int rc = flash_urdf_load(urdf_path, package_dirs, n_package_dirs, &u);
  if (rc) return rc;
  rc = flash_model_create(flash_urdf_desc(u), opt, out);
  flash_urdf_free(u);
  return rc;
}

int flash_read_point_cloud(const char* path, double origin[3], double* xyz, double* rgb, int64_t cap, int64_t* n) {
  if (!path || !n || cap < 0 || (cap > 0 && !xyz)) return fail(nullptr, FLASH_ERR_INVALID, "flash_read_point_cloud: bad arguments");
  std::vector<double> p, c; std::string err; double org[3] = {0, 0, 0};
  if (!flashurdf::read_point_cloud(path, p, rgb ? &c : nullptr, org, err)) return fail(nullptr, FLASH_ERR_IO, err);
  if (origin) { origin[0] = org[0]; origin[1] = org[1]; origin[2] = org[2]; }
  *n = (int64_t)(p.size() / 3);
  const size_t m = (size_t)std::min<int64_t>(cap, *n);
  if (m) std::memcpy(xyz, p.data(), m * 3 * sizeof(double));
  if (m && rgb) std::memcpy(rgb, c.data(), m * 3 * sizeof(double));
  return FLASH_OK;
}

int flash_num_positions(const flash_ctx* ctx) { return ctx ? ctx->hm.M.nq : FLASH_ERR_INVALID; }
int flash_num_states(const flash_ctx* ctx) { return ctx ? ctx->hm.M.nx : FLASH_ERR_INVALID; }
int flash_num_surfaces(const flash_ctx* ctx) { return ctx ? ctx->hm.M.n_surfaces : FLASH_ERR_INVALID; }
int64_t flash_num_points(const flash_ctx* ctx) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (ctx->subs.empty()) return ctx->ps[ctx->cur].n;
  int64_t n = 0;
  for (const flash_ctx* sub : ctx->subs) n += sub->ps[sub->cur].n;
  return n;
}

int flash_points_set(flash_ctx* ctx, const double* xyz, int64_t n) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (n < 0 || (n > 0 && !xyz)) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set: bad arguments");
  if (!ctx->subs.empty()) return multi_points_set<double>(ctx, xyz, n, false);
  return points_set_sync<double>(ctx, xyz, n);
}

int flash_points_set_f32(flash_ctx* ctx, const float* xyz, int64_t n) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (n < 0 || (n > 0 && !xyz)) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set_f32: bad arguments");
  if (!ctx->subs.empty()) return multi_points_set<float>(ctx, xyz, n, false);
  return points_set_sync<float>(ctx, xyz, n);
}

int flash_points_set_async(flash_ctx* ctx, const double* xyz, int64_t n) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (n < 0 || (n > 0 && !xyz)) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set_async: bad arguments");
  if (!ctx->subs.empty()) return multi_points_set<double>(ctx, xyz, n, true);
  const double* ptrs[1] = {xyz};
  const int64_t counts[1] = {n};
  int rc = ingest<double>(ctx, ctx->ps[ctx->cur ^ 1], ptrs, counts, 1); if (rc) return rc;
  ctx->back_pending = true;
  return FLASH_OK;
}

int flash_points_set_async_f32(flash_ctx* ctx, const float* xyz, int64_t n) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (n < 0 || (n > 0 && !xyz)) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set_async_f32: bad arguments");
  if (!ctx->subs.empty()) return multi_points_set<float>(ctx, xyz, n, true);
  const float* ptrs[1] = {xyz};
  const int64_t counts[1] = {n};
  int rc = ingest<float>(ctx, ctx->ps[ctx->cur ^ 1], ptrs, counts, 1); if (rc) return rc;
  ctx->back_pending = true;
  return FLASH_OK;
}

static int points_set_file_impl(flash_ctx* ctx, const char* path, double origin[3], bool async) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!path) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set_file: null path");
  std::vector<double> p; std::string err; double org[3] = {0, 0, 0};
  if (!flashurdf::read_point_cloud(path, p, nullptr, org, err)) return fail(ctx, FLASH_ERR_IO, err);
  if (origin) { origin[0] = org[0]; origin[1] = org[1]; origin[2] = org[2]; }
  if (!ctx->subs.empty()) return multi_points_set<double>(ctx, p.data(), (int64_t)(p.size() / 3), async);
  const double* ptrs[1] = {p.data()};
  const int64_t counts[1] = {(int64_t)(p.size() / 3)};
  // the parsed rows are pageable: ingest copies them into the ctx's pinned ring before it returns, the DMA reads the ring
  int rc = ingest<double>(ctx, ctx->ps[ctx->cur ^ 1], ptrs, counts, 1); if (rc) return rc;
  if (async) { ctx->back_pending = true; return FLASH_OK; }
  rc = swap_sets(ctx); if (rc) return rc;
  CU(cudaEventSynchronize(ctx->ps[ctx->cur].ready));
  return FLASH_OK;
}
int flash_points_set_file(flash_ctx* ctx, const char* path, double origin[3]) { return points_set_file_impl(ctx, path, origin, false); }
int flash_points_set_file_async(flash_ctx* ctx, const char* path, double origin[3]) { return points_set_file_impl(ctx, path, origin, true); }

int flash_points_swap(flash_ctx* ctx) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!ctx->subs.empty()) return multi_swap(ctx);
  if (!ctx->back_pending) return fail(ctx, FLASH_ERR_STATE, "flash_points_swap: no flash_points_set_async / flash_frames_set_async pending");
  CU(cudaSetDevice(ctx->device));
  return swap_sets(ctx);
}

int flash_points_sync(flash_ctx* ctx) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!ctx->subs.empty()) { for (flash_ctx* sub : ctx->subs) { int rc = flash_points_sync(sub); if (rc) { ctx->err = sub->err; return rc; } } return FLASH_OK; }
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->copy_stream));
  return FLASH_OK;
}

static int frames_check(flash_ctx* ctx, const void* const* xyz, const int64_t* counts, int32_t F, const char* who) {
  if (!xyz || !counts || F < 1 || F > MAX_FRAMES) return fail(ctx, FLASH_ERR_INVALID, std::string(who) + ": need 1.." + std::to_string(MAX_FRAMES) + " frames");
  long long tot = 0;
  for (int f = 0; f < F; ++f) { if (counts[f] < 0 || (counts[f] > 0 && !xyz[f])) return fail(ctx, FLASH_ERR_INVALID, std::string(who) + ": bad frame"); tot += counts[f]; }
  if (tot >= (1ll << 31)) return fail(ctx, FLASH_ERR_TOO_LARGE, std::string(who) + ": more than 2^31 points in one batch of frames");
  return 0;
}

int flash_frames_set(flash_ctx* ctx, const double* const* xyz, const int64_t* counts, int32_t F) {
  if (!ctx) return FLASH_ERR_INVALID;
  int rc = frames_check(ctx, (const void* const*)xyz, counts, F, "flash_frames_set"); if (rc) return rc;
  if (!ctx->subs.empty()) return multi_frames_set(ctx, xyz, counts, F, false);
  rc = ingest<double>(ctx, ctx->ps[ctx->cur ^ 1], xyz, counts, F); if (rc) return rc;
  rc = swap_sets(ctx); if (rc) return rc;
  CU(cudaEventSynchronize(ctx->ps[ctx->cur].ready));
  return FLASH_OK;
}

int flash_frames_set_async(flash_ctx* ctx, const double* const* xyz, const int64_t* counts, int32_t F) {
  if (!ctx) return FLASH_ERR_INVALID;
  int rc = frames_check(ctx, (const void* const*)xyz, counts, F, "flash_frames_set_async"); if (rc) return rc;
  if (!ctx->subs.empty()) return multi_frames_set(ctx, xyz, counts, F, true);
  rc = ingest<double>(ctx, ctx->ps[ctx->cur ^ 1], xyz, counts, F); if (rc) return rc;
  ctx->back_pending = true;
  return FLASH_OK;
}

int flash_num_frames(const flash_ctx* ctx) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!ctx->subs.empty()) return ctx->frame_sub.empty() ? (ctx->have_points ? 1 : 0) : (int)ctx->frame_sub.size();
  return ctx->ps[ctx->cur].set ? ctx->ps[ctx->cur].ft.n_frames : 0;
}

int flash_points_set_device(flash_ctx* ctx, const void* device_float4, int64_t n) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (n < 0 || (n > 0 && !device_float4)) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set_device: bad arguments");
  if (((uintptr_t)device_float4 & 15) != 0) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set_device: pointer must be 16-byte aligned");
  if (!ctx->subs.empty()) return fail(ctx, FLASH_ERR_INVALID, "flash_points_set_device: a multi-device context shards host points; give each device its own single-device context for device-resident clouds");
  CU(cudaSetDevice(ctx->device));
  flash_ctx::PointSet& P = ctx->ps[ctx->cur ^ 1];
  int rc = set_alloc(ctx, P, 32); if (rc) return rc;
  cudaStream_t cs = ctx->copy_stream;
  if (P.retired_valid) { CU(cudaStreamWaitEvent(cs, P.retired, 0)); P.retired_valid = false; }
  if (!P.external && P.d_pts) { CU(cudaStreamSynchronize(cs)); CU(cudaFree(P.d_pts)); }
  P.d_pts = (float4*)device_float4; P.cap_pts = 0; P.external = true; P.perm_on = false; P.grp_on = false;
  P.n = n; P.max_cnt = n; P.set = true;
  P.ft.n_frames = 1; P.ft.off[0] = 0; P.ft.off[1] = (n + 31) & ~31ll; P.ft.cnt[0] = n; P.ft.src_off[0] = 0; P.ft.src_off[1] = n;
  if (n > 0) {
    const long long groups = (n + 31) / 32;
    rc = ensure(ctx, P.d_grp, P.cap_grp, (size_t)groups); if (rc) return rc;
    groups_ball_kernel<<<(unsigned)((groups * 32 + 255) / 256), 256, 0, cs>>>(P.d_pts, n, P.d_grp);
    CU(cudaGetLastError());
    ctx->launches++;
    P.grp_on = true;
  }
  CU(cudaEventRecord(P.ready, cs));
  rc = swap_sets(ctx); if (rc) return rc;
  CU(cudaEventSynchronize(ctx->ps[ctx->cur].ready));
  return FLASH_OK;
}

// flash_cost_grad in stages, so that a multi-device parent can run them device by device around one reduction:
//   cg_begin   X -> device, prepare, group pre-pass, cost kernel (all enqueued, nothing waits)
//   (reduction: ncclAllReduce across processes, or the peer reduction across the parent's devices)
//   cg_end     epilogue + results into the pinned staging buffer (enqueued)
//   cg_finish  wait, hand the results to the caller
static int cg_begin(flash_ctx* ctx, const double* X, int32_t C, int cfg_per_frame, bool grad, bool want_argmin) {
  const ModelC& M = ctx->hm.M;
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  CU(cudaSetDevice(ctx->device));
  int rc = ensure_batch(ctx, C); if (rc) return rc;
  ctx->last = flash_timing{0, 0, 0, 0, 0, 0, 0};
  const size_t nX = (size_t)C * M.nx;
  std::memcpy(ctx->h_pin, X, nX * sizeof(double));
  rc = pull_pinned(ctx, ctx->d_X, ctx->h_pin, nX); if (rc) return rc;
  int32_t* d_arg = nullptr;
  if (want_argmin) {
    rc = ensure(ctx, ctx->d_argmin, ctx->cap_argmin, (size_t)C * (size_t)std::max<long long>(P.n, 1)); if (rc) return rc;
    d_arg = ctx->d_argmin;
  }
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[0], ctx->stream));
  if ((rc = run_prepare(ctx, C, true))) return rc;   // also zeroes the accumulators
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[1], ctx->stream));
  if ((rc = run_cost(ctx, C, grad, d_arg, cfg_per_frame))) return rc;
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[2], ctx->stream));
  return 0;
}
static int cg_end(flash_ctx* ctx, int32_t C, double w_def, bool grad) {
  const ModelC& M = ctx->hm.M;
  CU(cudaSetDevice(ctx->device));
  int rc;
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[3], ctx->stream));
  if ((rc = run_epilogue(ctx, C, w_def, grad))) return rc;
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[4], ctx->stream));
  const size_t nX = (size_t)C * M.nx;
  double* h_cost = ctx->h_pin + nX;
  double* h_grad = h_cost + C;
  CU(cudaMemcpyAsync(h_cost, ctx->d_cost, (size_t)C * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (grad) CU(cudaMemcpyAsync(h_grad, ctx->d_grad, nX * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  return 0;
}
static int cg_finish(flash_ctx* ctx, int32_t C, double* cost, double* grad) {
  const ModelC& M = ctx->hm.M;
  CU(cudaSetDevice(ctx->device));
  int rc = check_status(ctx); if (rc) return rc;   // synchronises the stream
  const size_t nX = (size_t)C * M.nx;
  if (cost) std::memcpy(cost, ctx->h_pin + nX, (size_t)C * sizeof(double));
  if (grad) std::memcpy(grad, ctx->h_pin + nX + C, nX * sizeof(double));
  return collect_timing(ctx);
}
static int cost_grad_impl(flash_ctx* ctx, const double* X, int32_t C, int cfg_per_frame, double w_def, double* cost, double* grad, int32_t* argmin, const char* who) {
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  int rc = cg_begin(ctx, X, C, cfg_per_frame, grad != nullptr, argmin != nullptr); if (rc) return rc;
  if ((rc = run_allreduce(ctx, C))) return rc;
  if ((rc = cg_end(ctx, C, w_def, grad != nullptr))) return rc;
  if (argmin && P.n > 0) CU(cudaMemcpyAsync(argmin, ctx->d_argmin, (size_t)C * P.n * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  (void)who;
  return cg_finish(ctx, C, cost, grad);
}


}  // extern "C" (templates below)

// ================================================================================================
// Single-process multi-GPU context (flash_options.n_devices > 1).  The reference's estimate_state is called from ONE
// Julia process (src/tracking.jl:8-27), so the library shards inside: points in round-robin chunks of the caller's
// order, frames round-robin, rays in contiguous ranges.  The accumulators (cost, wrenches, RBF adjoint sums: a few
// KB per configuration) are combined by ONE kernel on the first device that reads every device's buffer through
// NVLink peer memory in a fixed order and writes the sum back to all of them -- deterministic, no NCCL in this mode
// (the process-per-GPU mode keeps ncclAllReduce, flash_comm_init).
__global__ void reduce_peers_kernel(double* const* __restrict__ accs, int G, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double s = 0.0;
    for (int g = 0; g < G; ++g) s += accs[g][i];
    for (int g = 0; g < G; ++g) accs[g][i] = s;
  }
}
static int multi_fail(flash_ctx* P, flash_ctx* sub, int rc) { P->err = "device " + std::to_string(sub->device) + ": " + sub->err; return rc; }
#define MSUB(call) do { int rc_ = (call); if (rc_) return multi_fail(ctx, sub, rc_); } while (0)

// every sub's cost kernel has been enqueued: sum the accumulators over the devices
static int multi_reduce(flash_ctx* ctx, int C) {
  flash_ctx* s0 = ctx->subs[0];
  const int G = (int)ctx->subs.size();
  const size_t n = (size_t)C * s0->hm.M.acc_len;
  if (s0->peer_acc_C != s0->cap_C) {   // (re)built whenever ensure_batch reallocated the accumulators
    std::vector<double*> h(G);
    for (int g = 0; g < G; ++g) h[g] = ctx->subs[g]->d_acc;
    { flash_ctx* ctx = s0; CU(cudaSetDevice(s0->device)); if (!s0->d_peer_acc) CU(cudaMalloc((void**)&s0->d_peer_acc, 64 * sizeof(double*)));
      CU(cudaMemcpyAsync(s0->d_peer_acc, h.data(), G * sizeof(double*), cudaMemcpyHostToDevice, s0->stream)); CU(cudaStreamSynchronize(s0->stream)); }
    s0->peer_acc_C = s0->cap_C;
  }
  for (flash_ctx* sub : ctx->subs) { flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); CU(cudaEventRecord(sub->ev_cost, sub->stream)); }
  {
    flash_ctx* ctx = s0;
    CU(cudaSetDevice(s0->device));
    for (int g = 1; g < G; ++g) CU(cudaStreamWaitEvent(s0->stream, s0->parent->subs[g]->ev_cost, 0));
    reduce_peers_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 64), 256, 0, s0->stream>>>(s0->d_peer_acc, G, n);
    CU(cudaGetLastError());
    s0->launches++; s0->last.launches++;
    CU(cudaEventRecord(s0->ev_red, s0->stream));
  }
  for (int g = 1; g < G; ++g) { flash_ctx* ctx = s0->parent->subs[g]; CU(cudaSetDevice(ctx->device)); CU(cudaStreamWaitEvent(ctx->stream, s0->ev_red, 0)); }
  return 0;
}

// Points are dealt to the devices in chunks, round-robin: every device sees the whole scene (a contiguous split of a
// spatially ordered cloud gave one of two GPUs all the points near the robot: 14.6 ms against 11.9 ms for a shuffled
// cloud on K5), and a chunk is a contiguous piece of host memory, i.e. one DMA.  About 8 chunks per device, 1,024 to
// 65,536 points each.
template <class T>
static int multi_points_set(flash_ctx* ctx, const T* xyz, int64_t n, bool async) {
  const int G = (int)ctx->subs.size();
  long long chunk = (n / ((long long)G * 8) + 31) & ~31ll;
  chunk = std::min<long long>(std::max<long long>(chunk, 1024), 65536);
  ctx->seg_off.assign(G, {}); ctx->seg_len.assign(G, {});
  std::vector<std::vector<const T*>> ptrs(G);
  std::vector<std::vector<int64_t>> cnts(G);
  int g = 0;
  for (long long at = 0; at < n; at += chunk, g = (g + 1) % G) {
    const long long len = std::min<long long>(chunk, n - at);
    ctx->seg_off[g].push_back(at); ctx->seg_len[g].push_back(len);
    ptrs[g].push_back(xyz + 3 * at); cnts[g].push_back(len);
  }
  ctx->n_points = n; ctx->have_points = true;
  ctx->frame_sub.clear();
  for (g = 0; g < G; ++g) {
    flash_ctx* sub = ctx->subs[g];
    MSUB(ingest<T>(sub, sub->ps[sub->cur ^ 1], ptrs[g].data(), cnts[g].data(), (int)ptrs[g].size(), true));
    sub->back_pending = true;
  }
  if (async) { ctx->back_pending = true; return FLASH_OK; }
  for (flash_ctx* sub : ctx->subs) { MSUB(swap_sets(sub)); }
  for (flash_ctx* sub : ctx->subs) { flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); CU(cudaEventSynchronize(sub->ps[sub->cur].ready)); }
  return FLASH_OK;
}
static int multi_swap(flash_ctx* ctx) {
  if (!ctx->back_pending) return fail(ctx, FLASH_ERR_STATE, "flash_points_swap: no flash_points_set_async / flash_frames_set_async pending");
  for (flash_ctx* sub : ctx->subs) { flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); if (sub->back_pending) { int rc = swap_sets(sub); if (rc) return rc; } }
  ctx->back_pending = false;
  return FLASH_OK;
}
// frames round-robin over the devices (K4: 8 frames on 8 GPUs, one each); a device without a frame holds an empty set
static int multi_frames_set(flash_ctx* ctx, const double* const* xyz, const int64_t* counts, int32_t F, bool async) {
  const int G = (int)ctx->subs.size();
  ctx->frame_sub.assign(F, 0); ctx->frame_slot.assign(F, 0);
  std::vector<std::vector<const double*>> ptrs(G);
  std::vector<std::vector<int64_t>> cnts(G);
  for (int f = 0; f < F; ++f) { const int g = f % G; ctx->frame_sub[f] = g; ctx->frame_slot[f] = (int)ptrs[g].size(); ptrs[g].push_back(xyz[f]); cnts[g].push_back(counts[f]); }
  ctx->have_points = false;
  for (int g = 0; g < G; ++g) {
    flash_ctx* sub = ctx->subs[g];
    if (ptrs[g].empty()) { sub->back_pending = false; continue; }
    MSUB(ingest<double>(sub, sub->ps[sub->cur ^ 1], ptrs[g].data(), cnts[g].data(), (int)ptrs[g].size()));
    sub->back_pending = true;
  }
  if (async) { ctx->back_pending = true; return FLASH_OK; }
  for (flash_ctx* sub : ctx->subs) if (sub->back_pending) { MSUB(swap_sets(sub)); }
  for (flash_ctx* sub : ctx->subs) { flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); if (sub->ps[sub->cur].ready) CU(cudaEventSynchronize(sub->ps[sub->cur].ready)); }
  return FLASH_OK;
}
static void multi_timing(flash_ctx* ctx) {
  ctx->last = flash_timing{0, 0, 0, 0, 0, 0, 0};
  for (flash_ctx* sub : ctx->subs) {
    ctx->last.ms_prepare = std::max(ctx->last.ms_prepare, sub->last.ms_prepare); ctx->last.ms_main = std::max(ctx->last.ms_main, sub->last.ms_main);
    ctx->last.ms_allreduce = std::max(ctx->last.ms_allreduce, sub->last.ms_allreduce); ctx->last.ms_epilogue = std::max(ctx->last.ms_epilogue, sub->last.ms_epilogue);
    ctx->last.ms_total = std::max(ctx->last.ms_total, sub->last.ms_total); ctx->last.ms_group = std::max(ctx->last.ms_group, sub->last.ms_group);
    ctx->last.launches += sub->last.launches;
  }
}
static int multi_cost_grad(flash_ctx* ctx, const double* X, int32_t C, double w_def, double* cost, double* grad, int32_t* argmin) {
  if (!ctx->have_points) return fail(ctx, FLASH_ERR_STATE, "flash_cost_grad: call flash_points_set first (the ctx holds a batch of frames: use flash_cost_grad_frames)");
  const long long N = ctx->n_points;
  for (flash_ctx* sub : ctx->subs) { sub->profiling = ctx->profiling; MSUB(cg_begin(sub, X, C, 0, grad != nullptr, argmin != nullptr)); }
  int rc = multi_reduce(ctx, C); if (rc) return rc;
  for (size_t g = 0; g < ctx->subs.size(); ++g) {
    flash_ctx* sub = ctx->subs[g];
    MSUB(cg_end(sub, C, w_def, grad != nullptr));
    const long long n_g = sub->ps[sub->cur].n;
    if (argmin && n_g > 0) {   // this device's chunks: columns of argmin[C][N]
      flash_ctx* par = ctx; flash_ctx* ctx = sub;
      long long at = 0;
      for (size_t k = 0; k < par->seg_off[g].size(); ++k) {
        const long long len = par->seg_len[g][k];
        CU(cudaMemcpy2DAsync(argmin + par->seg_off[g][k], (size_t)N * sizeof(int32_t), sub->d_argmin + at, (size_t)n_g * sizeof(int32_t), (size_t)len * sizeof(int32_t), (size_t)C,
                             cudaMemcpyDeviceToHost, sub->stream));
        at += len;
      }
    }
  }
  for (size_t g = 0; g < ctx->subs.size(); ++g) { flash_ctx* sub = ctx->subs[g]; MSUB(cg_finish(sub, C, g == 0 ? cost : nullptr, g == 0 ? grad : nullptr)); }
  multi_timing(ctx);
  return FLASH_OK;
}
static int multi_cost_grad_frames(flash_ctx* ctx, const double* X, int32_t C, double w_def, double* cost, double* grad) {
  const int F = (int)ctx->frame_sub.size(), G = (int)ctx->subs.size();
  if (F == 0) return fail(ctx, FLASH_ERR_STATE, "flash_cost_grad_frames: call flash_frames_set first");
  const int nx = ctx->hm.M.nx;
  std::vector<std::vector<double>> Xg(G), cg(G), gg(G);
  std::vector<int> nf(G, 0);
  for (int f = 0; f < F; ++f) { const int g = ctx->frame_sub[f]; Xg[g].insert(Xg[g].end(), X + (size_t)f * C * nx, X + (size_t)(f + 1) * C * nx); nf[g]++; }
  for (int g = 0; g < G; ++g) {
    if (!nf[g]) continue;
    flash_ctx* sub = ctx->subs[g];
    sub->profiling = ctx->profiling;
    if ((long long)C * nf[g] > 65535 * 16) return fail(ctx, FLASH_ERR_TOO_LARGE, "flash_cost_grad_frames: frames x configs too large");
    MSUB(cg_begin(sub, Xg[g].data(), C * nf[g], C, grad != nullptr, false));
    MSUB(cg_end(sub, C * nf[g], w_def, grad != nullptr));      // frames are independent: no reduction
  }
  for (int g = 0; g < G; ++g) {
    if (!nf[g]) continue;
    flash_ctx* sub = ctx->subs[g];
    cg[g].resize((size_t)C * nf[g]); if (grad) gg[g].resize((size_t)C * nf[g] * nx);
    MSUB(cg_finish(sub, C * nf[g], cg[g].data(), grad ? gg[g].data() : nullptr));
  }
  for (int f = 0; f < F; ++f) {
    const int g = ctx->frame_sub[f], k = ctx->frame_slot[f];
    std::memcpy(cost + (size_t)f * C, cg[g].data() + (size_t)k * C, (size_t)C * sizeof(double));
    if (grad) std::memcpy(grad + (size_t)f * C * nx, gg[g].data() + (size_t)k * C * nx, (size_t)C * nx * sizeof(double));
  }
  multi_timing(ctx);
  return FLASH_OK;
}

static int multi_create(const flash_model_desc* desc, const flash_options* opt, flash_ctx** out) {
  const int G = opt->n_devices;
  if (G > 16) return fail(nullptr, FLASH_ERR_INVALID, "flash_options.n_devices > 16");
  flash_ctx* P = new flash_ctx;
  std::string err;
  if (!build_model(desc, opt, P->hm, err)) { delete P; return fail(nullptr, FLASH_ERR_INVALID, err); }
  P->device = opt->devices[0];
  for (int g = 0; g < G; ++g) {
    flash_options o = *opt;
    o.n_devices = 0; o.device = opt->devices[g];
    flash_ctx* sub = nullptr;
    const int rc = flash_model_create(desc, &o, &sub);
    if (rc) { const std::string msg = "device " + std::to_string(opt->devices[g]) + ": " + g_create_err; flash_destroy(P); return fail(nullptr, rc, msg); }
    sub->parent = P;
    P->subs.push_back(sub);
    if (cudaEventCreateWithFlags(&sub->ev_cost, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&sub->ev_red, cudaEventDisableTiming) != cudaSuccess) {
      flash_destroy(P); return fail(nullptr, FLASH_ERR_CUDA, "cudaEventCreate failed");
    }
  }
  // the reduction kernel runs on the first device and touches every other device's accumulators through peer memory
  flash_ctx* s0 = P->subs[0];
  cudaSetDevice(s0->device);
  for (int g = 1; g < G; ++g) {
    const int dev = P->subs[g]->device;
    if (dev == s0->device) continue;   // (the same device listed twice: legal, used by the single-GPU tests)
    int can = 0;
    cudaDeviceCanAccessPeer(&can, s0->device, dev);
    if (!can) { flash_destroy(P); return fail(nullptr, FLASH_ERR_CUDA, "device " + std::to_string(s0->device) + " cannot access device " + std::to_string(dev) + " as a peer (NVLink / PCIe P2P required)"); }
    const cudaError_t e = cudaDeviceEnablePeerAccess(dev, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { flash_destroy(P); return fail(nullptr, FLASH_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); }
    cudaGetLastError();
  }
  *out = P;
  return FLASH_OK;
}

extern "C" {

int flash_cost_grad(flash_ctx* ctx, const double* X, int32_t C, double w_def, double* cost, double* grad, int32_t* argmin) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!X || C < 1 || !cost) return fail(ctx, FLASH_ERR_INVALID, "flash_cost_grad: bad arguments");
  if (!ctx->subs.empty()) return multi_cost_grad(ctx, X, C, w_def, cost, grad, argmin);
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  if (!P.set) return fail(ctx, FLASH_ERR_STATE, "flash_cost_grad: call flash_points_set first");
  if (P.ft.n_frames != 1) return fail(ctx, FLASH_ERR_STATE, "flash_cost_grad: the ctx holds a batch of frames, use flash_cost_grad_frames");
  return cost_grad_impl(ctx, X, C, 0, w_def, cost, grad, argmin, "flash_cost_grad");
}

int flash_cost_grad_frames(flash_ctx* ctx, const double* X, int32_t C, double w_def, double* cost, double* grad) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!X || C < 1 || !cost) return fail(ctx, FLASH_ERR_INVALID, "flash_cost_grad_frames: bad arguments");
  if (!ctx->subs.empty()) return multi_cost_grad_frames(ctx, X, C, w_def, cost, grad);
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  if (!P.set) return fail(ctx, FLASH_ERR_STATE, "flash_cost_grad_frames: call flash_frames_set first");
  const long long tot = (long long)C * P.ft.n_frames;
  if (tot > 65535 * 16) return fail(ctx, FLASH_ERR_TOO_LARGE, "flash_cost_grad_frames: frames x configs too large");
  return cost_grad_impl(ctx, X, (int32_t)tot, C, w_def, cost, grad, nullptr, "flash_cost_grad_frames");
}

int flash_skin_eval(flash_ctx* ctx, const double* x, const double* xyz, int64_t n, double* dist, int32_t* argmin) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!x || n < 0 || (n > 0 && (!xyz || !dist))) return fail(ctx, FLASH_ERR_INVALID, "flash_skin_eval: bad arguments");
  if (!ctx->subs.empty()) { int rc = flash_skin_eval(ctx->subs[0], x, xyz, n, dist, argmin); if (rc) ctx->err = ctx->subs[0]->err; return rc; }
  if (n == 0) return FLASH_OK;
  const ModelC& M = ctx->hm.M;
  CU(cudaSetDevice(ctx->device));
  int rc = ensure_batch(ctx, 1); if (rc) return rc;
  ctx->last = flash_timing{0, 0, 0, 0, 0, 0, 0};
  CU(cudaMemcpyAsync(ctx->d_X, x, (size_t)M.nx * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = run_prepare(ctx, 1))) return rc;
  // query points: [n][3] doubles -> float4 (own scratch; the frame's points stay untouched)
  // scratch layout in doubles: xyz[3n] | pad to even | float4[n] (2 doubles each) | dist[n]
  const size_t off_f4 = ((size_t)n * 3 + 1) & ~(size_t)1, off_dist = off_f4 + 2 * (size_t)n;
  rc = ensure(ctx, ctx->d_qd, ctx->cap_qd, off_dist + (size_t)n); if (rc) return rc;
  rc = ensure(ctx, ctx->d_qi, ctx->cap_qi, (size_t)n); if (rc) return rc;
  double* d_xyz = ctx->d_qd; float4* d_f4 = (float4*)(ctx->d_qd + off_f4); double* d_dist = ctx->d_qd + off_dist;
  CU(cudaMemcpyAsync(d_xyz, xyz, (size_t)n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  convert_points_kernel<double><<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_xyz, d_f4, n);
  CU(cudaGetLastError());
  ctx->launches++; ctx->last.launches++;
  CostParams p;
  p.hot = ctx->d_hot; p.geo_f4 = ctx->d_geo_f4; p.geo_u32 = ctx->d_geo_u32; p.grid_pool = ctx->d_grid_pool;
  p.pts = d_f4; p.n_pts = n; p.tab = ctx->d_tab; p.acc = ctx->d_acc; p.argmin = ctx->d_qi; p.dist = d_dist; p.perm = nullptr; p.grp = nullptr; p.frames = nullptr; p.cfg_per_frame = 1; p.rec = nullptr; p.rec_stride = 0; p.part = nullptr; p.tile_order = nullptr;
  p.pts_per_block = choose_tile(ctx, n, 1); p.use_grid = ctx->use_grid ? 1 : 0;
  const int tiles = (int)((n + p.pts_per_block - 1) / p.pts_per_block);
  const size_t smem = cost_layout(ctx, false).total;
  if ((rc = launch_cost_t<false, true, true>(ctx, p, 1, tiles, smem))) return rc;
  CU(cudaMemcpyAsync(dist, d_dist, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (argmin) CU(cudaMemcpyAsync(argmin, ctx->d_qi, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  return check_status(ctx);
}

// ---- flash_skin_eval without any allocation (include/flashsdf.h: no counterpart in the reference) ----
// The scratch of the call: a file-scope device array, FLASH_STATIC_SCRATCH_BYTES on every GPU that loads the library.
//   [0, 256)                               status word of prepare_kernel (cleared by every call)
//   [256, FLASH_STATIC_STATE_BYTES)        x[nx] | tab[tab_len] | aux[aux_len]   (doubles, each rounded up to 256 B)
//   then, at fixed offsets sized for FLASH_STATIC_MAX_QUERY points: xyz (3 doubles) | float4 | dist (double) | argmin (int)
// In HOST code `g_static_scratch` is the host-side stand-in of the symbol, not a device address: the address a kernel may
// be given comes from cudaGetSymbolAddress, looked up once per device with that device current (static_scratch_base).
static __device__ __align__(256) unsigned char g_static_scratch[FLASH_STATIC_SCRATCH_BYTES];
static_assert((size_t)FLASH_STATIC_STATE_BYTES + (size_t)FLASH_STATIC_MAX_QUERY * (24 + 16 + 8 + 4) <= (size_t)FLASH_STATIC_SCRATCH_BYTES, "static scratch layout");
static std::mutex g_static_mu;                 // held for a whole call: the array is shared by every context of the process
static unsigned char* g_static_base[64] = {};  // per device ordinal

static int static_scratch_base(flash_ctx* ctx, unsigned char** base) {   // ctx->device is current; g_static_mu is held
  const int dev = ctx->device;
  if (dev >= 0 && dev < 64 && g_static_base[dev]) { *base = g_static_base[dev]; return 0; }
  void* p = nullptr;
  CU(cudaGetSymbolAddress(&p, g_static_scratch));
  if (!p) return fail(ctx, FLASH_ERR_CUDA, "cudaGetSymbolAddress(g_static_scratch) returned a null address");
  if (dev >= 0 && dev < 64) g_static_base[dev] = (unsigned char*)p;
  *base = (unsigned char*)p;
  return 0;
}

int flash_skin_eval_static(flash_ctx* ctx, const double* x, const double* xyz, int64_t n, double* dist, int32_t* argmin) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!x || n < 0 || (n > 0 && (!xyz || !dist))) return fail(ctx, FLASH_ERR_INVALID, "flash_skin_eval_static: bad arguments");
  if (!ctx->subs.empty()) { int rc = flash_skin_eval_static(ctx->subs[0], x, xyz, n, dist, argmin); if (rc) ctx->err = ctx->subs[0]->err; return rc; }
  if (n > FLASH_STATIC_MAX_QUERY)
    return fail(ctx, FLASH_ERR_TOO_LARGE, "flash_skin_eval_static: " + std::to_string((long long)n) + " query points exceed FLASH_STATIC_MAX_QUERY = " +
                                              std::to_string(FLASH_STATIC_MAX_QUERY) + " (split the call, or use flash_skin_eval)");
  const ModelC& M = ctx->hm.M;
  auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
  const size_t off_x = 256, off_tab = off_x + up((size_t)M.nx * 8), off_aux = off_tab + up((size_t)std::max(M.tab_len, 1) * 8),
               end_state = off_aux + up((size_t)std::max(M.aux_len, 1) * 8);
  if (end_state > (size_t)FLASH_STATIC_STATE_BYTES)
    return fail(ctx, FLASH_ERR_TOO_LARGE, "flash_skin_eval_static: the model's state and tables (" + std::to_string(end_state) +
                                              " bytes) exceed FLASH_STATIC_STATE_BYTES = " + std::to_string(FLASH_STATIC_STATE_BYTES));
  if (n == 0) return FLASH_OK;
  std::lock_guard<std::mutex> lock(g_static_mu);
  CU(cudaSetDevice(ctx->device));
  unsigned char* base = nullptr;
  int rc = static_scratch_base(ctx, &base); if (rc) return rc;
  const size_t Q = (size_t)FLASH_STATIC_MAX_QUERY, off_xyz = (size_t)FLASH_STATIC_STATE_BYTES, off_f4 = off_xyz + 24 * Q, off_dist = off_f4 + 16 * Q, off_arg = off_dist + 8 * Q;
  int* d_status = (int*)base;
  double *d_x = (double*)(base + off_x), *d_tab = (double*)(base + off_tab), *d_aux = (double*)(base + off_aux);
  double *d_xyz = (double*)(base + off_xyz), *d_dist = (double*)(base + off_dist);
  float4* d_f4 = (float4*)(base + off_f4);
  int32_t* d_arg = (int32_t*)(base + off_arg);
  ctx->last = flash_timing{0, 0, 0, 0, 0, 0, 0};
  CU(cudaMemsetAsync(d_status, 0, sizeof(int), ctx->stream));   // the array keeps the last call's contents: the flag must start at zero
  CU(cudaMemcpyAsync(d_x, x, (size_t)M.nx * sizeof(double), cudaMemcpyDefault, ctx->stream));
  CU(cudaMemcpyAsync(d_xyz, xyz, (size_t)n * 3 * sizeof(double), cudaMemcpyDefault, ctx->stream));
  const size_t psmem = prepare_smem_doubles(M) * sizeof(double);
  if (psmem > 48 * 1024) CU(cudaFuncSetAttribute(prepare_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psmem));
  prepare_kernel<<<1, PREP_T, psmem, ctx->stream>>>(ctx->d_model, d_x, d_tab, d_aux, nullptr, d_status);
  CU(cudaGetLastError());
  convert_points_kernel<double><<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_xyz, d_f4, n);
  CU(cudaGetLastError());
  ctx->launches += 2; ctx->last.launches += 2;
  CostParams p;
  p.hot = ctx->d_hot; p.geo_f4 = ctx->d_geo_f4; p.geo_u32 = ctx->d_geo_u32; p.grid_pool = ctx->d_grid_pool;
  p.pts = d_f4; p.n_pts = n; p.tab = d_tab; p.acc = nullptr; p.argmin = d_arg; p.dist = d_dist; p.perm = nullptr; p.grp = nullptr; p.frames = nullptr; p.cfg_per_frame = 1; p.rec = nullptr; p.rec_stride = 0; p.part = nullptr; p.tile_order = nullptr;
  p.pts_per_block = choose_tile(ctx, n, 1); p.use_grid = ctx->use_grid ? 1 : 0;
  const int tiles = (int)((n + p.pts_per_block - 1) / p.pts_per_block);
  if ((rc = launch_cost_t<false, true, true>(ctx, p, 1, tiles, cost_layout(ctx, false).total))) return rc;
  int st = 0;
  CU(cudaMemcpyAsync(dist, d_dist, (size_t)n * sizeof(double), cudaMemcpyDefault, ctx->stream));
  if (argmin) CU(cudaMemcpyAsync(argmin, d_arg, (size_t)n * sizeof(int32_t), cudaMemcpyDefault, ctx->stream));
  CU(cudaMemcpyAsync(&st, d_status, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (st != 0) return fail(ctx, FLASH_ERR_INVALID, "singular RBF system (coincident skin control points)");
  return FLASH_OK;
}

int flash_kinect_rays(int32_t rows, int32_t cols, double vertical_fov, double horizontal_fov, double* rays) {
  if (rows < 1 || cols < 1 || !rays) return FLASH_ERR_INVALID;
  kinect_rays(rows, cols, vertical_fov, horizontal_fov, rays);
  return FLASH_OK;
}

// ray casting in two halves so that a multi-device parent can start every device before it waits for any
static int raycast_launch(flash_ctx* ctx, const double* x, const double* rays, int64_t R, const double cam[12]) {
  const ModelC& M = ctx->hm.M;
  CU(cudaSetDevice(ctx->device));
  int rc = ensure_batch(ctx, 1); if (rc) return rc;
  ctx->last = flash_timing{0, 0, 0, 0, 0, 0, 0};
  CU(cudaMemcpyAsync(ctx->d_X, x, (size_t)M.nx * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = run_prepare(ctx, 1))) return rc;
  const int nblocks = (int)((R + BLOCK - 1) / BLOCK);
  // scratch: rays[3R] | depths[R] | hits[3R] | total(1)
  rc = ensure(ctx, ctx->d_qd, ctx->cap_qd, (size_t)R * 7 + 2); if (rc) return rc;
  rc = ensure(ctx, ctx->d_qi, ctx->cap_qi, (size_t)nblocks + 2); if (rc) return rc;
  double* d_rays = ctx->d_qd; double* d_depths = d_rays + 3 * R; double* d_hits = d_depths + R; long long* d_total = (long long*)(d_hits + 3 * R);
  CU(cudaMemcpyAsync(d_rays, rays, (size_t)R * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  RayParams p;
  p.hot = ctx->d_hot; p.geo_f4 = ctx->d_geo_f4; p.geo_u32 = ctx->d_geo_u32; p.grid_pool = ctx->d_grid_pool;
  p.tab = ctx->d_tab; p.rays = d_rays; p.R = R;
  for (int i = 0; i < 12; ++i) p.cam[i] = cam[i];
  p.depths = d_depths; p.block_hits = ctx->d_qi; p.use_grid = ctx->use_grid ? 1 : 0;
  const size_t smem = smem_layout(M.tab_len, 0, M.tabf_len, false).total;
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(raycast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[1], ctx->stream));
  raycast_kernel<<<nblocks, BLOCK, smem, ctx->stream>>>(p);
  CU(cudaGetLastError());
  if (ctx->profiling) CU(cudaEventRecord(ctx->ev[2], ctx->stream));
  scan_blocks_kernel<<<1, 1024, 0, ctx->stream>>>(ctx->d_qi, nblocks, d_total);
  CU(cudaGetLastError());
  compact_hits_kernel<<<nblocks, BLOCK, 0, ctx->stream>>>(d_rays, d_depths, R, ctx->d_qi, d_hits, nullptr, cam[0], cam[1], cam[2], cam[3], cam[4], cam[5], cam[6],
                                                          cam[7], cam[8], cam[9], cam[10], cam[11]);
  CU(cudaGetLastError());
  ctx->launches += 3; ctx->last.launches += 3;
  return 0;
}
static int raycast_collect(flash_ctx* ctx, int64_t R, double* depths, double* hits, int64_t* nhits) {
  CU(cudaSetDevice(ctx->device));
  double* d_rays = ctx->d_qd; double* d_depths = d_rays + 3 * R; double* d_hits = d_depths + R; long long* d_total = (long long*)(d_hits + 3 * R);
  long long total = 0;
  CU(cudaMemcpyAsync(depths, d_depths, (size_t)R * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(&total, d_total, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
  int rc = check_status(ctx); if (rc) return rc;
  if (hits && total > 0) { CU(cudaMemcpyAsync(hits, d_hits, (size_t)total * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream)); CU(cudaStreamSynchronize(ctx->stream)); }
  if (nhits) *nhits = total;
  if (ctx->profiling) { CU(cudaEventElapsedTime(&ctx->last.ms_main, ctx->ev[1], ctx->ev[2])); ctx->last.ms_total = ctx->last.ms_main; }
  return FLASH_OK;
}

int flash_raycast(flash_ctx* ctx, const double* x, const double* rays, int64_t R, const double cam[12], double* depths, double* hits, int64_t* nhits) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!x || R < 0 || !cam || (R > 0 && (!rays || !depths))) return fail(ctx, FLASH_ERR_INVALID, "flash_raycast: bad arguments");
  if (nhits) *nhits = 0;
  if (R == 0) return FLASH_OK;
  if (!ctx->subs.empty()) {
    // rays sharded in contiguous ranges (image row bands for a row-major sensor): depths land at their ray's index, hits are
    // concatenated in device order = ray order (raycast_points, src/depthsensors.jl:99-113)
    const int G = (int)ctx->subs.size();
    std::vector<int64_t> off(G + 1);
    for (int g = 0; g <= G; ++g) off[g] = (int64_t)((__int128)R * g / G);
    for (int g = 0; g < G; ++g) {
      flash_ctx* sub = ctx->subs[g];
      sub->profiling = ctx->profiling;
      if (off[g + 1] > off[g]) MSUB(raycast_launch(sub, x, rays + 3 * off[g], off[g + 1] - off[g], cam));
    }
    int64_t total = 0;
    for (int g = 0; g < G; ++g) {
      flash_ctx* sub = ctx->subs[g];
      if (off[g + 1] == off[g]) continue;
      int64_t nh = 0;
      MSUB(raycast_collect(sub, off[g + 1] - off[g], depths + off[g], hits ? hits + 3 * total : nullptr, &nh));
      total += nh;
    }
    if (nhits) *nhits = total;
    multi_timing(ctx);
    return FLASH_OK;
  }
  int rc = raycast_launch(ctx, x, rays, R, cam); if (rc) return rc;
  return raycast_collect(ctx, R, depths, hits, nhits);
}

static int run_iterate(flash_ctx* ctx, int C, const flash_solver_params* sp, double n_total, bool have_div, size_t it_smem) {
  iterate_kernel<<<C, PREP_T, it_smem, ctx->stream>>>(ctx->d_model, ctx->d_X, ctx->d_tab, ctx->d_aux, ctx->d_acc, C, sp->deformation_cost_weight, ctx->d_cost,
                                                  ctx->d_grad, n_total > 0 ? 1.0 / n_total : 0.0, sp->rate, sp->max_step, sp->gradient_convergence_tolerance,
                                                  have_div ? ctx->d_div : nullptr, ctx->d_done, ctx->d_iters, ctx->d_status);
  CU(cudaGetLastError());
  ctx->launches++; ctx->last.launches++;
  return 0;
}
// state upload of flash_estimate_state (one device)
static int es_begin(flash_ctx* ctx, const double* X, int32_t C, const double* divisors, size_t it_smem) {
  const ModelC& M = ctx->hm.M;
  CU(cudaSetDevice(ctx->device));
  int rc = ensure_batch(ctx, C); if (rc) return rc;
  ctx->last = flash_timing{0, 0, 0, 0, 0, 0, 0};
  const size_t nX = (size_t)C * M.nx;
  std::memcpy(ctx->h_pin, X, nX * sizeof(double));
  rc = pull_pinned(ctx, ctx->d_X, ctx->h_pin, nX); if (rc) return rc;
  CU(cudaMemsetAsync(ctx->d_done, 0, (size_t)C * sizeof(int), ctx->stream));
  CU(cudaMemsetAsync(ctx->d_iters, 0, (size_t)C * sizeof(int), ctx->stream));
  if (divisors) {
    if (!ctx->d_div) CU(cudaMalloc((void**)&ctx->d_div, (size_t)M.nx * sizeof(double)));
    double* h_div = ctx->h_pin + (size_t)C * (2 * (size_t)M.nx + 2);
    std::memcpy(h_div, divisors, (size_t)M.nx * sizeof(double));
    rc = pull_pinned(ctx, ctx->d_div, h_div, (size_t)M.nx); if (rc) return rc;
  }
  if (it_smem > 48 * 1024) CU(cudaFuncSetAttribute(iterate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)it_smem));
  return 0;
}
// The solver loop as ONE persistent kernel (track_kernel): one configuration, no communicator; models with hulls only on
// frame-sized clouds.  Anything else runs the two-kernel loop below.
static bool track_kernel_applies(const flash_ctx* ctx, int C) {
  static const bool off = [] { const char* e = getenv("FLASHSDF_NO_TRACK_KERNEL"); return e && e[0] == '1'; }();
  const ModelC& M = ctx->hm.M;
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  if (off || C != 1 || ctx->comm || P.n <= 0) return false;
  // models with hulls: the persistent loop runs without group records (every warp culls the hulls for its own 32 points),
  // which beats group_kernel + cost_kernel + two launches per iteration on a depth frame (IRB140, 307,200 points, 30
  // iterations: 1.64 ms against 1.90 ms; the K4 scene: 2.51 against 2.62) but not on a large cloud (K5: +22 % per pass)
  if (P.grp_on && ctx->group_cull && M.n_convex >= 2 && P.n > 524288) return false;
  return track_layout(M).total + 64 <= (size_t)ctx->smem_optin;
}
static int run_track(flash_ctx* ctx, const flash_solver_params* sp, double n_total, bool have_div, bool want_cost) {
  const ModelC& M = ctx->hm.M;
  const flash_ctx::PointSet& P = ctx->ps[ctx->cur];
  long long nblk = std::min<long long>(ctx->sm_count, (P.n + 31) / 32);
  long long ppb = (((P.n + nblk - 1) / nblk + 31) / 32) * 32;
  nblk = (P.n + ppb - 1) / ppb;
  int rc = ensure(ctx, ctx->d_part, ctx->cap_part, std::max<size_t>((size_t)2 * (size_t)nblk, 3) * (size_t)M.acc_len); if (rc) return rc;
  if (!ctx->deterministic) CU(cudaMemsetAsync(ctx->d_part, 0, (size_t)3 * M.acc_len * sizeof(double), ctx->stream));
  if (!ctx->d_bar) CU(cudaMalloc((void**)&ctx->d_bar, 64 * sizeof(unsigned)));
  CU(cudaMemsetAsync(ctx->d_bar, 0, 64 * sizeof(unsigned), ctx->stream));
  TrackParams t;
  CostParams& p = t.cp;
  p.hot = ctx->d_hot; p.geo_f4 = ctx->d_geo_f4; p.geo_u32 = ctx->d_geo_u32; p.grid_pool = ctx->d_grid_pool;
  p.pts = P.d_pts; p.n_pts = P.n; p.tab = nullptr; p.acc = nullptr; p.argmin = nullptr; p.dist = nullptr; p.perm = nullptr; p.grp = nullptr;
  p.rec = nullptr; p.rec_stride = 0; p.frames = nullptr; p.cfg_per_frame = 1; p.pts_per_block = (int)ppb; p.use_grid = ctx->use_grid ? 1 : 0; p.part = nullptr; p.tile_order = nullptr;
  t.model = ctx->d_model; t.X = ctx->d_X; t.part = ctx->d_part; t.bar = ctx->d_bar; t.cost = ctx->d_cost; t.iters = ctx->d_iters; t.done = ctx->d_done; t.status = ctx->d_status;
  t.divisors = have_div ? ctx->d_div : nullptr;
  t.w_def = sp->deformation_cost_weight; t.inv_n = n_total > 0 ? 1.0 / n_total : 0.0; t.rate = sp->rate; t.max_step = sp->max_step; t.tol = sp->gradient_convergence_tolerance;
  t.iteration_limit = sp->iteration_limit; t.want_cost = want_cost ? 1 : 0;
  const size_t smem = track_layout(M).total;
  const void* k = ctx->deterministic ? (const void*)track_kernel<true, true> : (ppb <= 32 * CWARPS ? (const void*)track_kernel<false, true> : (const void*)track_kernel<false, false>);
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  void* args[] = {&t};
  CU(cudaLaunchCooperativeKernel(k, dim3((unsigned)nblk), dim3(CBLOCK), args, smem, ctx->stream));
  ctx->launches++; ctx->last.launches++;
  return 0;
}
// The solver loop on every device of a multi-device parent: each device evaluates its shard, the peer reduction hands
// all of them the same sums, and every device then takes the same NaiveSolver step on its own copy of the state.
static int multi_estimate_state(flash_ctx* ctx, double* X, int32_t C, const flash_solver_params* sp, const double* divisors, double* cost_out, int32_t* iters_out) {
  if (!ctx->have_points) return fail(ctx, FLASH_ERR_STATE, "flash_estimate_state: call flash_points_set first");
  const ModelC& M = ctx->hm.M;
  const size_t it_smem = std::max(prepare_smem_doubles(M), epilogue_smem_doubles(M)) * sizeof(double);
  const double n_total = (double)ctx->n_points;
  for (flash_ctx* sub : ctx->subs) { sub->profiling = false; MSUB(es_begin(sub, X, C, divisors, it_smem)); MSUB(run_prepare(sub, C, true)); }
  for (int it = 0; it < sp->iteration_limit; ++it) {
    for (flash_ctx* sub : ctx->subs) { flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); int rc = run_cost(sub, C, true, nullptr); if (rc) return rc; }
    int rc = multi_reduce(ctx, C); if (rc) return rc;
    for (flash_ctx* sub : ctx->subs) { flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); rc = run_iterate(sub, C, sp, n_total, divisors != nullptr, it_smem); if (rc) return rc; }
  }
  if (cost_out) {
    for (flash_ctx* sub : ctx->subs) { flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); int rc = run_cost(sub, C, false, nullptr); if (rc) return rc; }
    int rc = multi_reduce(ctx, C); if (rc) return rc;
    { flash_ctx* sub = ctx->subs[0]; flash_ctx* ctx = sub; CU(cudaSetDevice(sub->device)); rc = run_epilogue(sub, C, sp->deformation_cost_weight, false); if (rc) return rc; }
  }
  const size_t nX = (size_t)C * M.nx;
  for (size_t g = 0; g < ctx->subs.size(); ++g) {
    flash_ctx* sub = ctx->subs[g];
    flash_ctx* ctx = sub;
    CU(cudaSetDevice(sub->device));
    if (g == 0) {
      CU(cudaMemcpyAsync(sub->h_pin, sub->d_X, nX * sizeof(double), cudaMemcpyDeviceToHost, sub->stream));
      if (cost_out) CU(cudaMemcpyAsync(sub->h_pin + nX, sub->d_cost, (size_t)C * sizeof(double), cudaMemcpyDeviceToHost, sub->stream));
      if (iters_out) CU(cudaMemcpyAsync(iters_out, sub->d_iters, (size_t)C * sizeof(int), cudaMemcpyDeviceToHost, sub->stream));
    }
    int rc = check_status(sub); if (rc) return rc;
  }
  std::memcpy(X, ctx->subs[0]->h_pin, nX * sizeof(double));
  if (cost_out) std::memcpy(cost_out, ctx->subs[0]->h_pin + nX, (size_t)C * sizeof(double));
  multi_timing(ctx);
  return FLASH_OK;
}

int flash_estimate_state(flash_ctx* ctx, double* X, int32_t C, const flash_solver_params* sp, const double* divisors, double* cost_out, int32_t* iters_out) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!X || C < 1 || !sp || sp->iteration_limit < 0) return fail(ctx, FLASH_ERR_INVALID, "flash_estimate_state: bad arguments");
  if (!ctx->subs.empty()) return multi_estimate_state(ctx, X, C, sp, divisors, cost_out, iters_out);
  if (!ctx->ps[ctx->cur].set) return fail(ctx, FLASH_ERR_STATE, "flash_estimate_state: call flash_points_set first");
  if (ctx->ps[ctx->cur].ft.n_frames != 1) return fail(ctx, FLASH_ERR_STATE, "flash_estimate_state: the ctx holds a batch of frames");
  const ModelC& M = ctx->hm.M;
  const size_t it_smem = std::max(prepare_smem_doubles(M), epilogue_smem_doubles(M)) * sizeof(double);
  const size_t nX = (size_t)C * M.nx;
  int rc = es_begin(ctx, X, C, divisors, it_smem); if (rc) return rc;
  // total point count across ranks (the reference divides by length(sensed_points), src/tracking.jl:20)
  double n_total = (double)ctx->ps[ctx->cur].n;
  if (ctx->comm && ctx->n_global >= 0) n_total = ctx->n_global;   // reduced once per point set, not once per call
  else if (ctx->comm) {
    double* d_n = ctx->d_cost;  // scratch
    CU(cudaMemcpyAsync(d_n, &n_total, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    int nrc = g_nccl.AllReduce(d_n, d_n, 1, NCCL_FLOAT64, NCCL_SUM, ctx->comm, ctx->stream);
    if (nrc != 0) return fail(ctx, FLASH_ERR_NCCL, "ncclAllReduce(point count) failed");
    CU(cudaMemcpyAsync(&n_total, d_n, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->n_global = n_total;
  }
  const bool prof = ctx->profiling; ctx->profiling = false;
  cudaEvent_t e0 = ctx->ev[0], e1 = ctx->ev[4];
  if (prof) CU(cudaEventRecord(e0, ctx->stream));
  // per iteration: cost_kernel -> [allreduce] -> iterate_kernel (epilogue + NaiveSolver step + prepare of the next pass)
  if (track_kernel_applies(ctx, C)) {
    if ((rc = run_track(ctx, sp, n_total, divisors != nullptr, cost_out != nullptr))) { ctx->profiling = prof; return rc; }
  } else {
    if ((rc = run_prepare(ctx, C, true))) { ctx->profiling = prof; return rc; }
    for (int it = 0; it < sp->iteration_limit; ++it) {
      if ((rc = run_cost(ctx, C, true, nullptr)) || (rc = run_allreduce(ctx, C))) { ctx->profiling = prof; return rc; }
      if ((rc = run_iterate(ctx, C, sp, n_total, divisors != nullptr, it_smem))) { ctx->profiling = prof; return rc; }
    }
    if (cost_out) {   // cost at the final state: the last iterate_kernel already prepared it
      if ((rc = run_cost(ctx, C, false, nullptr)) || (rc = run_allreduce(ctx, C)) || (rc = run_epilogue(ctx, C, sp->deformation_cost_weight, false))) { ctx->profiling = prof; return rc; }
    }
  }
  if (prof) CU(cudaEventRecord(e1, ctx->stream));
  ctx->profiling = prof;
  double* h_cost = ctx->h_pin + nX;
  CU(cudaMemcpyAsync(ctx->h_pin, ctx->d_X, nX * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (cost_out) CU(cudaMemcpyAsync(h_cost, ctx->d_cost, (size_t)C * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (iters_out) CU(cudaMemcpyAsync(iters_out, ctx->d_iters, (size_t)C * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  rc = check_status(ctx); if (rc) return rc;
  std::memcpy(X, ctx->h_pin, nX * sizeof(double));
  if (cost_out) std::memcpy(cost_out, h_cost, (size_t)C * sizeof(double));
  if (prof) CU(cudaEventElapsedTime(&ctx->last.ms_total, e0, e1));
  return FLASH_OK;
}

int flash_contour_mesh(flash_ctx* ctx, const double* x, int32_t surface, const double lo[3], const double hi[3], double spacing, double iso,
                       float* triangles, int64_t cap, int64_t* n_triangles) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!ctx->subs.empty()) { int rc = flash_contour_mesh(ctx->subs[0], x, surface, lo, hi, spacing, iso, triangles, cap, n_triangles); if (rc) ctx->err = ctx->subs[0]->err; return rc; }
  const ModelC& M = ctx->hm.M;
  if (!x || !lo || !hi || !n_triangles || !(spacing > 0) || cap < 0 || surface >= M.n_surfaces)
    return fail(ctx, FLASH_ERR_INVALID, "flash_contour_mesh: bad arguments");
  *n_triangles = 0;
  GridParams g;
  long long nodes = 1;
  for (int k = 0; k < 3; ++k) {
    if (!(hi[k] > lo[k])) return fail(ctx, FLASH_ERR_INVALID, "flash_contour_mesh: empty region");
    g.org[k] = lo[k];
    const double cells = std::ceil((hi[k] - lo[k]) / spacing);
    if (!(cells < 4096)) return fail(ctx, FLASH_ERR_TOO_LARGE, "flash_contour_mesh: more than 4096 cells per axis; increase spacing");
    g.dim[k] = (int)cells + 1;
    nodes *= g.dim[k];
  }
  if (nodes > (1ll << 30)) return fail(ctx, FLASH_ERR_TOO_LARGE, "flash_contour_mesh: more than 2^30 grid nodes; increase spacing");
  CU(cudaSetDevice(ctx->device));
  int rc = ensure_batch(ctx, 1); if (rc) return rc;
  ctx->last = flash_timing{0, 0, 0, 0, 0, 0, 0};
  CU(cudaMemcpyAsync(ctx->d_X, x, (size_t)M.nx * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  if ((rc = run_prepare(ctx, 1))) return rc;
  // scratch (in doubles): field[nodes] floats | counter | triangles[cap][9] floats
  const size_t off_cnt = ((size_t)nodes + 1) / 2, off_tri = off_cnt + 1;
  rc = ensure(ctx, ctx->d_qd, ctx->cap_qd, off_tri + ((size_t)cap * 9 + 1) / 2 + 1); if (rc) return rc;
  float* d_field = (float*)ctx->d_qd;
  unsigned long long* d_cnt = (unsigned long long*)(ctx->d_qd + off_cnt);
  float* d_tri = (float*)(ctx->d_qd + off_tri);
  g.hot = ctx->d_hot; g.geo_f4 = ctx->d_geo_f4; g.geo_u32 = ctx->d_geo_u32;
  g.tab = ctx->d_tab; g.h = spacing; g.surface = surface; g.field = d_field;
  const size_t smem = smem_layout(M.tab_len, 0, M.tabf_len, false).total;
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(field_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  field_grid_kernel<<<(unsigned)((nodes + BLOCK - 1) / BLOCK), BLOCK, smem, ctx->stream>>>(g);
  CU(cudaGetLastError());
  CU(cudaMemsetAsync(d_cnt, 0, sizeof(unsigned long long), ctx->stream));
  MarchParams mp;
  mp.field = d_field; mp.h = spacing; mp.iso = (float)iso; mp.tris = (triangles && cap > 0) ? d_tri : nullptr; mp.cap = (unsigned long long)cap; mp.count = d_cnt;
  for (int k = 0; k < 3; ++k) { mp.org[k] = lo[k]; mp.dim[k] = g.dim[k]; }
  const long long cells = (long long)(g.dim[0] - 1) * (g.dim[1] - 1) * (g.dim[2] - 1);
  march_tets_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, ctx->stream>>>(mp);
  CU(cudaGetLastError());
  ctx->launches += 2; ctx->last.launches += 2;
  unsigned long long cnt = 0;
  CU(cudaMemcpyAsync(&cnt, d_cnt, sizeof cnt, cudaMemcpyDeviceToHost, ctx->stream));
  rc = check_status(ctx); if (rc) return rc;   // synchronises
  *n_triangles = (int64_t)cnt;
  const unsigned long long take = std::min<unsigned long long>(cnt, (unsigned long long)cap);
  if (triangles && take > 0) CU(cudaMemcpy(triangles, d_tri, (size_t)take * 9 * sizeof(float), cudaMemcpyDeviceToHost));
  return FLASH_OK;
}

// ---- multi-GPU
int flash_comm_unique_id(void* id128) {
  std::string err;
  if (!id128) return FLASH_ERR_INVALID;
  if (!load_nccl(err)) { g_create_err = err; return FLASH_ERR_NCCL; }
  nccl_uid id;
  int rc = g_nccl.GetUniqueId(&id);
  if (rc != 0) { g_create_err = "ncclGetUniqueId failed"; return FLASH_ERR_NCCL; }
  std::memcpy(id128, &id, sizeof id);
  return FLASH_OK;
}

int flash_comm_init(flash_ctx* ctx, int32_t rank, int32_t world, const void* id128) {
  if (ctx && !ctx->subs.empty()) return fail(ctx, FLASH_ERR_INVALID, "flash_comm_init: a multi-device context already spans its devices");
  if (!ctx) return FLASH_ERR_INVALID;
  if (world < 1 || rank < 0 || rank >= world || !id128) return fail(ctx, FLASH_ERR_INVALID, "flash_comm_init: bad arguments");
  std::string err;
  if (!load_nccl(err)) return fail(ctx, FLASH_ERR_NCCL, err);
  CU(cudaSetDevice(ctx->device));
  nccl_uid id; std::memcpy(&id, id128, sizeof id);
  if (ctx->comm) { g_nccl.CommDestroy(ctx->comm); ctx->comm = nullptr; }
  int rc = g_nccl.CommInitRank(&ctx->comm, world, id, rank);
  if (rc != 0) { ctx->comm = nullptr; return fail(ctx, FLASH_ERR_NCCL, std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error")); }
  ctx->rank = rank; ctx->world = world;
  return FLASH_OK;
}

int flash_comm_world(const flash_ctx* ctx) { return ctx ? ctx->world : FLASH_ERR_INVALID; }

// ---- measurement
int flash_profiling(flash_ctx* ctx, int32_t enable) { if (!ctx) return FLASH_ERR_INVALID; ctx->profiling = enable != 0; for (flash_ctx* sub : ctx->subs) sub->profiling = enable != 0; return FLASH_OK; }
int flash_last_timing(const flash_ctx* ctx, flash_timing* t) { if (!ctx || !t) return FLASH_ERR_INVALID; *t = ctx->last; return FLASH_OK; }
int64_t flash_launch_count(const flash_ctx* ctx) {
  if (!ctx) return FLASH_ERR_INVALID;
  int64_t n = ctx->launches;
  for (const flash_ctx* sub : ctx->subs) n += sub->launches;
  return n;
}
int flash_num_devices(const flash_ctx* ctx) { return ctx ? (ctx->subs.empty() ? 1 : (int)ctx->subs.size()) : FLASH_ERR_INVALID; }
int flash_device_info(const flash_ctx* ctx, int32_t* sm_count, int32_t* clock_khz, int64_t* l2_bytes) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!ctx->subs.empty()) return flash_device_info(ctx->subs[0], sm_count, clock_khz, l2_bytes);
  if (sm_count) *sm_count = ctx->sm_count;
  if (clock_khz) *clock_khz = ctx->clock_khz;
  if (l2_bytes) *l2_bytes = ctx->l2_bytes;
  return FLASH_OK;
}
int flash_grid_report(const flash_ctx* ctx, int32_t surface, int32_t dims[3], double* spacing, double* max_abs_err) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!ctx->subs.empty()) return flash_grid_report(ctx->subs[0], surface, dims, spacing, max_abs_err);
  const ModelC& M = ctx->hm.M;
  if (surface < 0 || surface >= M.n_surfaces || M.surf_kind[surface] != SURF_CONVEX || !ctx->use_grid) return FLASH_ERR_INVALID;
  const ConvexC& C = M.cvx[M.surf_slot[surface]];
  if (dims) for (int k = 0; k < 3; ++k) dims[k] = C.g_dim[k];
  if (spacing) *spacing = 1.0 / C.g_inv_h;
  if (max_abs_err) *max_abs_err = ctx->grid_err[M.surf_slot[surface]];
  return FLASH_OK;
}

int flash_grid_download(const flash_ctx* ctx, int32_t surface, double origin[3], double geometry_to_body[12], float* values) {
  if (!ctx) return FLASH_ERR_INVALID;
  if (!ctx->subs.empty()) return flash_grid_download(ctx->subs[0], surface, origin, geometry_to_body, values);
  const ModelC& M = ctx->hm.M;
  if (surface < 0 || surface >= M.n_surfaces || M.surf_kind[surface] != SURF_CONVEX || !ctx->use_grid) return FLASH_ERR_INVALID;
  const ConvexC& C = M.cvx[M.surf_slot[surface]];
  if (origin) for (int k = 0; k < 3; ++k) origin[k] = C.g_org[k];
  if (geometry_to_body) for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) geometry_to_body[4 * i + j] = C.frame[3 * i + j]; geometry_to_body[4 * i + 3] = C.frame[9 + i]; }
  if (values) {
    const size_t n = (size_t)C.g_dim[0] * C.g_dim[1] * C.g_dim[2];
    if (cudaSetDevice(ctx->device) != cudaSuccess) return FLASH_ERR_CUDA;
    if (cudaMemcpy(values, ctx->d_grid_pool + C.g_off, n * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess) return FLASH_ERR_CUDA;
  }
  return FLASH_OK;
}

}  // extern "C"
