// hostcheck.cpp -- TEST INFRASTRUCTURE.  Compiles the product's __host__ __device__ math
// (csrc/flash_dev.h, csrc/flash_host.h) with g++ and replays, serially, exactly what the CUDA
// kernels do per config and per point, so the formulas can be checked against the oracle on a box
// without a GPU.  Never linked into libflashsdf.so; the product has no CPU path.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../point-cloud-signed-distance_b200/csrc/flash_host.h"

using namespace flashdev;
using namespace flashhost;

struct HC { HostModel hm; HotModel hot; };
static int g_group_cull = 1;            // replay the warp-level hull culling of cost_kernel
static int g_skin_cull = 1;             // ... and the group-level skin cull
static int g_exact = 1;                 // ... with exact centre distances and BSP start nodes (group_prepass)
static int g_dbg = 0;
static long long g_skin_groups_skipped = 0;
static long long g_cand_total = 0, g_groups = 0, g_rounds_or = 0, g_rounds_max = 0;
static unsigned g_or = 0; static int g_max = 0;
static thread_local std::string g_err;

extern "C" {

const char* hc_last_error() { return g_err.c_str(); }

void* hc_model_create(const flash_model_desc* d, const flash_options* o) {
  HC* h = new HC;
  if (!build_model(d, o, h->hm, g_err)) { delete h; return nullptr; }
  make_hot(h->hm.M, h->hot);
  return h;
}
void hc_model_destroy(void* p) { delete (HC*)p; }

// out: nx, nq, tab_len, acc_len, aux_len, n_convex, n_skins
void hc_info(void* p, int* out) {
  ModelC& M = ((HC*)p)->hm.M;
  out[0] = M.nx; out[1] = M.nq; out[2] = M.tab_len; out[3] = M.acc_len; out[4] = M.aux_len; out[5] = M.n_convex; out[6] = M.n_skins;
}
void hc_convex_info(void* p, int j, int* nverts, int* nplanes, float* lo, float* hi, float* depth_bound) {
  ConvexC& C = ((HC*)p)->hm.M.cvx[j];
  *nverts = C.nverts; *nplanes = C.nplanes; *depth_bound = C.depth_bound;
  for (int k = 0; k < 3; ++k) { lo[k] = C.lo[k]; hi[k] = C.hi[k]; }
}

// prepare kernel, one config
void hc_prepare(void* p, const double* x, double* tab, double* aux) {
  ModelC& M = ((HC*)p)->hm.M;
  fk_config(M, x, aux);
  prepare_convex(M, aux, tab);
  for (int k = 0; k < M.n_skins; ++k) {
    const SkinC& K = M.skin[k];
    skin_centres(M, K, x, aux, tab);
    double* LU = aux + aux_skin0(M) + K.aux_off;
    skin_assemble(M, K, tab, LU, 0, 1);
    lu_factor(LU, LU + K.M * K.M, K.M, 0, 1);
    skin_solve(M, K, LU, LU + K.M * K.M, tab);
  }
}

// main kernel, one config, serial over points; fp32 contributions for convex winners (as the warp
// shuffles carry them), fp64 for skins, fp64 accumulation
void hc_accumulate(void* p, const double* tab, const float* pts, long long n, int use_grid, double* acc, double* dist, int* argmin) {
  HC* h = (HC*)p;
  const HotModel& H = h->hot;
  GeoView G; G.f4 = h->hm.geo_f4.data(); G.u32 = h->hm.geo_u32.data(); G.grid_pool = nullptr;
  std::vector<float> tabf_v(H.tabf_len + 4);
  float* tabf = tabf_v.data();
  tabf_fill(H, tab, tabf, 0, 1);
  const unsigned all_convex = H.n_convex >= 32 ? 0xffffffffu : ((1u << H.n_convex) - 1u);
  unsigned cand = all_convex; int first = -1;
  unsigned skin_skip = 0u; float cut0 = 3.0e38f; unsigned long long starts = 0ull;
  for (long long i = 0; i < n; ++i) {
    if (g_group_cull && (i & 31) == 0 && H.n_convex > 1) {
      // the warp's 32 points: bounding ball as flash_ingest.cuh warp_ball computes it (Badoiu-Clarkson from the centroid)
      const int cntp = (int)std::min<long long>(32, n - i);
      float cx = 0.f, cy = 0.f, cz = 0.f;
      for (int k = 0; k < cntp; ++k) { cx += pts[3 * (i + k)]; cy += pts[3 * (i + k) + 1]; cz += pts[3 * (i + k) + 2]; }
      cx /= cntp; cy /= cntp; cz /= cntp;
      float mm = 0.f;
      for (int it = 0; it <= 10; ++it) {
        int who = 0; mm = -1.f;
        for (int k = 0; k < cntp; ++k) {
          const float dx = pts[3 * (i + k)] - cx, dy = pts[3 * (i + k) + 1] - cy, dz = pts[3 * (i + k) + 2] - cz;
          const float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
          if (d2 > mm) { mm = d2; who = k; }
        }
        if (it == 10) break;
        const float sgl = 1.0f / (float)(it + 2);
        cx = fmaf(pts[3 * (i + who)] - cx, sgl, cx); cy = fmaf(pts[3 * (i + who) + 1] - cy, sgl, cy); cz = fmaf(pts[3 * (i + who) + 2] - cz, sgl, cz);
      }
      const float rho = sqrtf(fmaxf(mm, 0.f)) * (1.f + 2e-6f) + 1e-30f;
      const GroupPre gp = group_prepass(H, G, tabf, cx, cy, cz, rho, g_exact != 0 && !use_grid, g_skin_cull != 0);
      cand = gp.cand; first = gp.first; skin_skip = gp.skip; cut0 = gp.cut; starts = gp.starts;
      g_cand_total += __builtin_popcount(cand); g_groups += 1;
      g_skin_groups_skipped += skin_skip ? 1 : 0;
      if (g_dbg) printf("group %lld first %d starts %llx cand %x rho %g c %g %g %g\n", i / 32, first, starts, cand, rho, cx, cy, cz);
    }
    PointEval e;
#ifdef FLASH_COUNTERS
    counters().exact_mask = 0u;
#endif
    eval_point(H, G, tab, tabf, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], cand, first, use_grid != 0, true, e, skin_skip, cut0, starts);
#ifdef FLASH_COUNTERS
    // exact narrow-phase rounds a warp of 32 consecutive points executes: one per hull ANY lane evaluates (the
    // kernel's candidate loop is warp-uniform) vs the most any single lane needs
    g_or |= counters().exact_mask; g_max = std::max(g_max, __builtin_popcount(counters().exact_mask));
    if ((i & 31) == 31 || i == n - 1) { g_rounds_or += __builtin_popcount(g_or); g_rounds_max += g_max; g_or = 0; g_max = 0; }
#endif
    double dv = e.is_skin ? e.sk.s : (double)e.d;
    if (dist) dist[i] = dv;
    if (argmin) argmin[i] = e.arg;
    if (!acc) continue;
    acc[0] += dv * dv;
    if (!e.is_skin) {
      float g = -2.f * e.d;
      float v0 = g * e.n.x, v1 = g * e.n.y, v2 = g * e.n.z;
      float t0 = e.y.y * v2 - e.y.z * v1, t1 = e.y.z * v0 - e.y.x * v2, t2 = e.y.x * v1 - e.y.y * v0;
      double* a = acc + 1 + 6 * e.slot;
      a[0] += v0; a[1] += v1; a[2] += v2; a[3] += t0; a[4] += t1; a[5] += t2;
    } else {
      const SkinC& K = H.skin[e.slot];
      const double* c = tab + K.tab_off;
      double* gY = acc + K.acc_off; double* gC = gY + K.npts + 4;
      for (int ci = 0; ci <= K.npts; ++ci) {
        double o[4];
        skin_adjoint_term(c, c + 3 * K.npts, K.npts, H.normalize != 0, ci, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], e.sk.s, e.sk.ginv, e.sk.nh, o);
        if (ci < K.npts) { gY[ci] += o[0]; gC[3 * ci] += o[1]; gC[3 * ci + 1] += o[2]; gC[3 * ci + 2] += o[3]; }
        else { gY[ci] += o[0]; gY[ci + 1] += o[1]; gY[ci + 2] += o[2]; gY[ci + 3] += o[3]; }
      }
    }
  }
}

void hc_epilogue(void* p, const double* x, const double* tab, const double* aux, const double* acc, double w_def, double* cost, double* grad) {
  ModelC& M = ((HC*)p)->hm.M;
  std::vector<double> scratch(6 * MAX_BODIES + MAX_SKIN_M);
  epilogue_config(M, x, tab, aux, acc, w_def, scratch.data(), cost, grad);
}

void hc_set_group_cull(int on) { g_group_cull = on; }
void hc_set_skin_cull(int on) { g_skin_cull = on; }
void hc_set_exact(int on) { g_exact = on; }
void hc_set_dbg(int on) { g_dbg = on; }
long long hc_skin_groups_skipped(int reset) { long long v = g_skin_groups_skipped; if (reset) g_skin_groups_skipped = 0; return v; }
// certified lower bound of skin k over a ball (tab = prepared fp64 table of one config)
float hc_skin_ball_lower(void* p, const double* tab, int k, float cx, float cy, float cz, float rho) {
  HC* h = (HC*)p;
  const HotModel& H = h->hot;
  std::vector<float> tabf(H.tabf_len + 4);
  tabf_fill(H, tab, tabf.data(), 0, 1);
  const SkinC& K = H.skin[k];
  return skin_ball_lower_f32(tabf.data() + K.tabf_off, K.npts, H.normalize != 0, cx, cy, cz, rho);
}
// out: candidate hulls summed over groups, groups
void hc_group_stats(long long* out, int reset) { out[0] = g_cand_total; out[1] = g_groups; if (reset) { g_cand_total = 0; g_groups = 0; } }
void hc_round_stats(long long* out, int reset) { out[0] = g_rounds_or; out[1] = g_rounds_max; if (reset) { g_rounds_or = 0; g_rounds_max = 0; } }

// instrumentation (only meaningful when built with -DFLASH_COUNTERS)
void hc_counters(long long* out, int reset) {
#ifdef FLASH_COUNTERS
  Counters& c = counters();
  out[0] = c.points; out[1] = c.narrow; out[2] = c.lb_tests; out[3] = c.skin64; out[4] = c.bsp_levels;
  if (reset) c = Counters{0, 0, 0, 0, 0, 0u};
#else
  for (int i = 0; i < 5; ++i) out[i] = -1;
#endif
}

// certified fp32 interval of skin slot k's value (skin_bound_f32) next to the fp64 value (skin_eval)
void hc_skin_bound(void* p, int k, const double* tab, const float* pts, long long n, float* lo, float* hi, double* s64) {
  HC* h = (HC*)p;
  const HotModel& H = h->hot;
  const SkinC& K = H.skin[k];
  std::vector<float> tabf(H.tabf_len + 4);
  tabf_fill(H, tab, tabf.data(), 0, 1);
  const double* c = tab + K.tab_off;
  for (long long i = 0; i < n; ++i) {
    SkinBound b = skin_bound_f32(tabf.data() + K.tabf_off, K.npts, H.normalize != 0, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
    lo[i] = b.lo; hi[i] = b.hi;
    s64[i] = skin_eval(c, c + 3 * K.npts, K.npts, H.normalize != 0, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]).s;
  }
}

// signed distance of geometry-frame points y[n][3] to convex slot j
void hc_hull_sd(void* p, int j, const float* y, long long n, float* d, float* nrm) {
  HC* h = (HC*)p;
  const ConvexC& C = h->hm.M.cvx[j];
  GeoView G; G.f4 = h->hm.geo_f4.data(); G.u32 = h->hm.geo_u32.data(); G.grid_pool = nullptr;
  for (long long i = 0; i < n; ++i) {
    F3 nn;
    d[i] = hull_sd(hull_view(G, C), f3(y[3 * i], y[3 * i + 1], y[3 * i + 2]), nn);
    nrm[3 * i] = nn.x; nrm[3 * i + 1] = nn.y; nrm[3 * i + 2] = nn.z;
  }
}

// the same through the closest-feature BSP tree (flash_bsp.h / hull_sd_bsp); leaf[] receives the leaf words
int hc_hull_sd_bsp(void* p, int j, const float* y, long long n, float* d, float* nrm, int* leaf) {
  HC* h = (HC*)p;
  const ConvexC& C = h->hm.M.cvx[j];
  GeoView G; G.f4 = h->hm.geo_f4.data(); G.u32 = h->hm.geo_u32.data(); G.grid_pool = nullptr;
  const HullView V = hull_view(G, C);
  for (long long i = 0; i < n; ++i) {
    F3 nn; const F3 yy = f3(y[3 * i], y[3 * i + 1], y[3 * i + 2]);
    const unsigned k = bsp_locate(V, 0u, yy);
    d[i] = bsp_feature_sd(V, k, yy, nn);
    nrm[3 * i] = nn.x; nrm[3 * i + 1] = nn.y; nrm[3 * i + 2] = nn.z;
    if (leaf) leaf[i] = (int)k;
  }
  return 1;
}
// out[5]: nodes, max depth, mean leaf depth, cells, dropped slivers
void hc_bsp_stats(void* p, int j, double* out) {
  HC* h = (HC*)p;
  for (int k = 0; k < 5; ++k) out[k] = j < (int)h->hm.bsp_stats.size() ? h->hm.bsp_stats[j][k] : 0.0;
}
// depth of the leaf each query reaches
void hc_bsp_depth(void* p, int j, const float* y, long long n, int* depth) {
  HC* h = (HC*)p;
  const ConvexC& C = h->hm.M.cvx[j];
  const F4* nodes = h->hm.geo_f4.data() + C.off_bn; const uint32_t* kids = h->hm.geo_u32.data() + C.off_bk;
  for (long long i = 0; i < n; ++i) {
    unsigned k = 0u; int dep = 0;
    do {
      const F4 q = nodes[k]; const uint32_t w = kids[k];
      const float s = fmaf(q.x, y[3 * i], fmaf(q.y, y[3 * i + 1], fmaf(q.z, y[3 * i + 2], q.w)));
      k = s > 0.f ? (w >> 16) : (w & 0xffffu); ++dep;
    } while (!(k & 0x8000u));
    depth[i] = dep;
  }
}

}  // extern "C"

