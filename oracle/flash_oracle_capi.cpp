// flash_oracle_capi.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see flash_oracle.hpp header).
// ctypes-facing C entry points of the CPU fp64 oracle.  The model description is the same
// flash_model_desc the product library takes, so tests feed both from one Python builder.
#include <atomic>
#include <cstdio>
#include <thread>

#include "flash_oracle.hpp"

using namespace fo;

namespace {

thread_local std::string g_err;

// dynamic-schedule parallel loop over [0, n) on `threads` std::threads (no OpenMP dependency)
template <class F>
void parallel_for(int64_t n, int threads, int64_t grain, F&& fn) {
  if (threads <= 1 || n <= grain) { for (int64_t i = 0; i < n; ++i) fn(i); return; }
  std::atomic<int64_t> next(0);
  std::string err;
  std::atomic<bool> failed(false);
  auto work = [&]() {
    try {
      for (;;) {
        int64_t b = next.fetch_add(grain);
        if (b >= n) break;
        for (int64_t i = b; i < std::min(n, b + grain); ++i) fn(i);
      }
    } catch (const std::exception& e) {
      if (!failed.exchange(true)) err = e.what();
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  if (failed) throw std::runtime_error(err);
}

// cost with the per-point loop spread over `threads` OpenMP threads; block partial sums are
// combined in block order, so the result does not depend on the thread count.
template <class T>
T cost_mt(const Model& m, const T* x, const double* pts, int64_t n, double w_def, int mode, int32_t* argmin, int threads) {
  PosedSurfaces<T> P = pose_surfaces<T>(m, x, mode);
  const int64_t B = 256;
  const int64_t nb = (n + B - 1) / B;
  std::vector<T> parts((size_t)nb);
  parallel_for(nb, threads, 8, [&](int64_t b) {
    T part = T(0.0);
    for (int64_t i = b * B; i < std::min(n, (b + 1) * B); ++i) {
      int a;
      T d = skin_eval<T>(P, V3d{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]}, &a);
      if (argmin) argmin[i] = a;
      part = part + d * d;
    }
    parts[(size_t)b] = part;
  });
  T c = T(0.0);
  for (int64_t b = 0; b < nb; ++b) c = c + parts[(size_t)b];
  T reg = T(0.0);
  for (int k = m.nq; k < m.nx(); ++k) reg = reg + x[k] * x[k];
  return c + w_def * reg;
}

// One ForwardDiff chunk: seeds x[k0 .. k0+N) and returns the partials (ForwardDiff
// chunk_mode_gradient!, traceback at examples/irb_and_squishable.ipynb:479-483).
template <int N>
double grad_chunk(const Model& m, const double* x, int k0, const double* pts, int64_t n, double w_def, int mode, int threads, double* g) {
  const int nx = m.nx();
  std::vector<Dual<N>> xd(nx);
  for (int k = 0; k < nx; ++k) xd[k] = Dual<N>(x[k]);
  for (int j = 0; j < N && k0 + j < nx; ++j) xd[k0 + j].d[j] = 1.0;
  Dual<N> c = cost_mt<Dual<N>>(m, xd.data(), pts, n, w_def, mode, nullptr, threads);
  for (int j = 0; j < N && k0 + j < nx; ++j) g[k0 + j] = c.d[j];
  return c.v;
}

double grad_dual(const Model& m, const double* x, const double* pts, int64_t n, double w_def, int mode, int threads, double* g, int* passes) {
  const int nx = m.nx();
  const int N = pick_chunk(nx);
  double c = 0;
  int np = 0;
  for (int k0 = 0; k0 < nx; k0 += N, ++np) {
    switch (N) {
      case 1: c = grad_chunk<1>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 2: c = grad_chunk<2>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 3: c = grad_chunk<3>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 4: c = grad_chunk<4>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 5: c = grad_chunk<5>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 6: c = grad_chunk<6>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 7: c = grad_chunk<7>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 8: c = grad_chunk<8>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      case 9: c = grad_chunk<9>(m, x, k0, pts, n, w_def, mode, threads, g); break;
      default: c = grad_chunk<10>(m, x, k0, pts, n, w_def, mode, threads, g); break;
    }
  }
  if (passes) *passes = np;
  return c;
}

}  // namespace

#define FO_TRY try {
#define FO_CATCH                                    \
  }                                                 \
  catch (const std::exception& e) {                 \
    g_err = e.what();                               \
    return -1;                                      \
  }

extern "C" {

const char* fo_last_error() { return g_err.c_str(); }

void* fo_model_create(const flash_model_desc* d, int normalize_by_gradient, int poly_degree) {
  try {
    return new Model(model_from_desc(d, normalize_by_gradient, poly_degree));
  } catch (const std::exception& e) {
    g_err = e.what();
    return nullptr;
  }
}
void fo_model_destroy(void* h) { delete (Model*)h; }
int fo_num_states(void* h) { return ((Model*)h)->nx(); }
int fo_num_positions(void* h) { return ((Model*)h)->nq; }
int fo_max_threads() { unsigned n = std::thread::hardware_concurrency(); return n ? (int)n : 1; }

int fo_hull_info(void* h, int s, int* nverts, int* nplanes, int* ntris) {
  Model& m = *(Model*)h;
  if (s < 0 || s >= (int)m.surfaces.size() || m.surfaces[s].kind != FLASH_SURF_CONVEX) return -1;
  *nverts = (int)m.surfaces[s].hull.verts.size();
  *nplanes = (int)m.surfaces[s].hull.planes.size();
  *ntris = (int)m.surfaces[s].hull.tris.size();
  return 0;
}

// (f::CostFunctor)(x), src/gradientdescent.jl:49-57 -- raw cost (+ argmin per point)
int fo_cost(void* h, const double* x, const double* pts, int64_t n, double w_def, int convex_mode, int threads, double* cost, int32_t* argmin) {
  FO_TRY
  *cost = cost_mt<double>(*(Model*)h, x, pts, n, w_def, convex_mode, argmin, threads);
  return 0;
  FO_CATCH
}

// ForwardDiff.gradient!(out, cost, x) in chunk mode (src/gradientdescent.jl:59-65)
int fo_cost_grad_dual(void* h, const double* x, const double* pts, int64_t n, double w_def, int convex_mode, int threads, double* cost, double* grad, int* passes) {
  FO_TRY
  *cost = grad_dual(*(Model*)h, x, pts, n, w_def, convex_mode, threads, grad, passes);
  return 0;
  FO_CATCH
}

// single-pass analytic gradient (see "analytic (adjoint) gradient" in flash_oracle.hpp); brute-force
// closest feature for convex bodies regardless of convex_mode's GJK setting for the value
int fo_cost_grad_adjoint(void* h, const double* x, const double* pts, int64_t n, double w_def, int convex_mode, int threads, double* cost, double* grad) {
  FO_TRY
  const Model& m = *(Model*)h;
  AdjointCtx A = adjoint_prepare(m, x, convex_mode);
  const int64_t B = 256, nb = (n + B - 1) / B;
  if (threads < 1) threads = 1;
  std::vector<AdjointAcc> accs((size_t)nb);
  parallel_for(nb, threads, 4, [&](int64_t b) {
    AdjointAcc& a = accs[(size_t)b]; a.init(m);
    for (int64_t i = b * B; i < std::min(n, (b + 1) * B); ++i) adjoint_point(A, V3d{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]}, a, nullptr);
  });
  AdjointAcc tot; tot.init(m);
  for (auto& a : accs) tot.add(a);
  adjoint_finish(A, tot, x, w_def, grad);
  *cost = tot.cost;
  return 0;
  FO_CATCH
}

// Flash.skin(state)(x), src/Flash.jl:265-268
int fo_skin_eval(void* h, const double* x, const double* pts, int64_t n, int convex_mode, int threads, double* dist, int32_t* argmin) {
  FO_TRY
  const Model& m = *(Model*)h;
  PosedSurfaces<double> P = pose_surfaces<double>(m, x, convex_mode);
  parallel_for(n, threads, 64, [&](int64_t i) {
    int a;
    dist[i] = skin_eval<double>(P, V3d{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]}, &a);
    if (argmin) argmin[i] = a;
  });
  return 0;
  FO_CATCH
}

// per-surface values (no min): out[n][S]
int fo_surface_eval(void* h, const double* x, const double* pts, int64_t n, int convex_mode, double* out) {
  FO_TRY
  const Model& m = *(Model*)h;
  PosedSurfaces<double> P = pose_surfaces<double>(m, x, convex_mode);
  const size_t S = m.surfaces.size();
  for (int64_t i = 0; i < n; ++i) {
    V3d p{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]};
    for (size_t s = 0; s < S; ++s)
      out[i * S + s] = m.surfaces[s].kind == FLASH_SURF_CONVEX ? convex_sd<double>(m.surfaces[s].hull, P.poses[s], p, convex_mode) : P.skins[s].eval(p);
  }
  return 0;
  FO_CATCH
}

int fo_kinect_rays(int rows, int cols, double vfov, double hfov, double* rays) {
  kinect_rays(rows, cols, vfov, hfov, rays);
  return 0;
}

// raycast_depths + raycast_points, src/depthsensors.jl:88-113
int fo_raycast(void* h, const double* x, const double* rays, int64_t R, const double* cam, int convex_mode, int threads, double* depths, double* hits, int64_t* nhits) {
  FO_TRY
  const Model& m = *(Model*)h;
  PosedSurfaces<double> P = pose_surfaces<double>(m, x, convex_mode);
  Xf<double> T = xf_from12<double>(cam);
  V3d origin = T.t;  // sensor_tform(0)
  parallel_for(R, threads, 16, [&](int64_t i) {
    V3d ray = T.rot(V3d{rays[3 * i], rays[3 * i + 1], rays[3 * i + 2]});  // rays_in_world (:83-86)
    ray = scale(ray, 1.0 / norm(ray));                                      // normalize(rays[i]) (:93)
    depths[i] = do_raycast(origin, ray, [&](const V3d& p) { return skin_eval<double>(P, p, nullptr); });
  });
  int64_t k = 0;
  for (int64_t i = 0; i < R; ++i)
    if (!std::isnan(depths[i])) {
      if (hits) {
        V3d r{rays[3 * i], rays[3 * i + 1], rays[3 * i + 2]};
        r = scale(r, depths[i] / norm(r));  // sensor_tform(dist * normalize(sensor.rays[row, col])) (:108)
        V3d w = T.apply(r);
        hits[3 * k] = w.x; hits[3 * k + 1] = w.y; hits[3 * k + 2] = w.z;
      }
      ++k;
    }
  if (nhits) *nhits = k;
  return 0;
  FO_CATCH
}

// estimate_state with a NaiveSolver (src/tracking.jl:8-27).  The step rule of
// SimpleGradientDescent @0fcc1f9 is not in the reference tree; restated from the keyword names
// (rate, max_step, iteration_limit, gradient_convergence_tolerance, precondition_divisors):
//   g = grad(cost/N); stop if |g| < tol; step = -rate*g./div; |step| clamped to max_step; x += step.
int fo_estimate_state(void* h, double* x, const double* pts, int64_t n, const flash_solver_params* sp, const double* divisors, int convex_mode, int threads, double* cost_out, int* iters_out, double* trace /*[iteration_limit][1+nx] or NULL*/) {
  FO_TRY
  const Model& m = *(Model*)h;
  const int nx = m.nx();
  std::vector<double> g(nx);
  int it = 0;
  for (; it < sp->iteration_limit; ++it) {
    double c = grad_dual(m, x, pts, n, sp->deformation_cost_weight, convex_mode, threads, g.data(), nullptr);
    if (trace) { trace[(size_t)it * (1 + nx)] = c; std::memcpy(trace + (size_t)it * (1 + nx) + 1, x, sizeof(double) * nx); }
    double gn = 0;
    for (int k = 0; k < nx; ++k) { g[k] /= (double)n; gn += g[k] * g[k]; }
    if (sp->gradient_convergence_tolerance > 0 && std::sqrt(gn) < sp->gradient_convergence_tolerance) break;
    double sn = 0;
    for (int k = 0; k < nx; ++k) { g[k] = -sp->rate * g[k] / (divisors ? divisors[k] : 1.0); sn += g[k] * g[k]; }
    sn = std::sqrt(sn);
    double sc = sn > sp->max_step ? sp->max_step / sn : 1.0;
    for (int k = 0; k < nx; ++k) x[k] += sc * g[k];
  }
  if (cost_out) *cost_out = cost_mt<double>(m, x, pts, n, sp->deformation_cost_weight, convex_mode, nullptr, threads);
  if (iters_out) *iters_out = it;
  return 0;
  FO_CATCH
}

}  // extern "C"
