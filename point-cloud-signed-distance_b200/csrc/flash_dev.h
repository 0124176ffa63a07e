// flash_dev.h -- device-side data layout and per-point / per-config math of libflashsdf.
//
// Everything here is `__host__ __device__`: the CUDA kernels in flash_kernels.cuh are the only
// product callers; the functions also compile with plain g++ so that tests/hostcheck can run the
// SAME arithmetic on a CPU-only box against the oracle before GPU time is spent (test
// infrastructure only -- the shipped library has no CPU path).
//
// Reference rows: A3 state layout, A4 quaternion normalisation, A5 cost, A6 min over
// surfaces, A7 RBF skin, A8 convex signed distance.  Citations are /root/reference paths.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define FLASH_HD __host__ __device__ __forceinline__
#define FLASH_HDN __host__ __device__ __noinline__
#else
#define FLASH_HD inline
#define FLASH_HDN inline
#endif

namespace flashdev {

#if defined(FLASH_COUNTERS) && !defined(__CUDA_ARCH__)
struct Counters { long long points, narrow, inside, gjk_iters, plane_tests, vert_tests, lb_tests, skin64, feat_evals, warm; unsigned exact_mask; };   // exact_mask: hulls whose narrow phase ran for the LAST point
inline Counters& counters() { static Counters c = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0u}; return c; }
#define FLASH_COUNT(field, n) (flashdev::counters().field += (n))
#else
#define FLASH_COUNT(field, n) ((void)0)
#endif

constexpr int MAX_BODIES = 32;
constexpr int MAX_SURFACES = 32;
constexpr int MAX_CONVEX = 24;
constexpr int MAX_SKINS = 4;
constexpr int MAX_SKIN_PTS = 64;           // control points per skin (two_link_arm has 46)
constexpr int MAX_SKIN_M = MAX_SKIN_PTS + 4;
constexpr int MAX_HULL_VERTS = 255;        // vertex ids are packed in 8 bits
constexpr int GJK_MAX_ITER = 24;

enum { JOINT_FIXED = 0, JOINT_REVOLUTE = 1, JOINT_QUAT = 2 };
enum { SURF_CONVEX = 0, SURF_RIGID_SKIN = 1, SURF_DEFORMABLE_SKIN = 2 };

struct F3 { float x, y, z; };
struct alignas(16) F4 { float x, y, z, w; };
struct D3 { double x, y, z; };
struct XfD { double R[9]; double t[3]; };  // p_out = R p + t

struct BodyC {
  int parent, jtype, qoff, pad;
  double axis[3];    // unit, joint frame
  double j2p[12];    // row-major [R|t], frameBefore -> parent body
};

// per convex surface; the geometry frame is the surface frame re-centred on the hull centroid
struct ConvexC {
  int surface, body, nverts, nplanes;
  int off_planes, off_tris, off_verts, off_vz;  // indices into the geometry blob (units: F4 / u32 / F4 / F4)
  int nplanes_pad, nverts_pad;               // counts rounded up to a multiple of 8 (padding is inert)
  int off_voff, off_vlist;                   // support direction map (u32 indices into the blob), see dir_cell()
  int off_pgrp, pgrp_pad;                    // plane-group hint map (u32 index into the blob; u16 per direction cell), see hull_plane_hint()
  float lo[3], hi[3];                        // AABB of the vertices, geometry frame
  float depth_bound, r_in;                   // >= max penetration depth; <= inradius about the geometry origin
  double frame[12];                          // geometry frame -> body frame
  // grid mode (body-frame SDF grid): origin, spacing, dims, offset into the grid pool (floats)
  float g_org[3]; float g_inv_h; int g_dim[3]; int g_on; long long g_off;
  // closest-feature BSP tree (flash_bsp.h): node planes / plain plane table / vertices / edges (F4 indices), child words (u32 index)
  int bsp_on, off_bn, off_bk, off_bp, off_bv, off_be;
};

struct SkinC {
  int surface, kind, npts, nsurf;
  int deform_off;  // first deformation scalar (relative to x + nq), -1 if rigid
  int M;           // npts + (poly_deg >= 1 ? 4 : 1)
  int tab_off;     // config table: centres[3n], w[n], a0, a[3]      (doubles)
  int acc_off;     // accumulators: gY[n+4], gC[3n]                  (doubles)
  int aux_off;     // per-config aux: LU[M*M], piv[M]                (doubles)
  int pt_off;      // into ModelC::skin_pt
  int tabf_off;    // fp32 table (floats): {cx,cy,cz,3w}[n], {a0,ax,ay,az}   (certified pre-filter)
  int pad;
};

struct ModelC {
  int n_bodies, n_surfaces, n_convex, n_skins;
  int nq, nx, normalize, poly_deg;
  int tab_len;   // doubles per config:   12*n_convex + sum(4n+4)
  int acc_len;   // doubles per config:   1 + 6*n_convex + sum(4n+4)
  int aux_len;   // doubles per config:   24*n_bodies + n_bodies + sum(M*M + M)
  int geo_f4, geo_u32;  // blob sizes
  int tabf_len;  // floats per config of the fp32 side table: 12*n_convex (padded to 4) + sum(4n+4)
  BodyC body[MAX_BODIES];
  ConvexC cvx[MAX_CONVEX];
  SkinC skin[MAX_SKINS];
  int surf_kind[MAX_SURFACES];
  int surf_slot[MAX_SURFACES];  // index into cvx[] or skin[]
  double skin_pt[MAX_SKINS * MAX_SKIN_PTS][3];
  int skin_pt_body[MAX_SKINS * MAX_SKIN_PTS];
};

// what the per-point code reads; copied to shared memory by every block of the main kernel
struct HotModel {
  int n_convex, n_skins, normalize, acc_len;
  int tab_len, tabf_len, pad[2];
  SkinC skin[MAX_SKINS];
  ConvexC cvx[MAX_CONVEX];
};

// offsets inside the per-config aux block
FLASH_HD int aux_X(const ModelC& M, int b) { return 12 * b; }
FLASH_HD int aux_P(const ModelC& M, int b) { return 12 * M.n_bodies + 12 * b; }
FLASH_HD int aux_qnorm(const ModelC& M, int b) { return 24 * M.n_bodies + b; }
FLASH_HD int aux_skin0(const ModelC& M) { return 25 * M.n_bodies; }

// unroll factor of the per-centre skin loops (the cost kernel is instruction-cache sensitive: flash_kernels.cuh CBLOCK)
#define FLASH_PRAGMA_(x) _Pragma(#x)
#define FLASH_PRAGMA(x) FLASH_PRAGMA_(x)
#ifndef FLASH_SKIN_UNROLL
#define FLASH_SKIN_UNROLL 2   // K5: 1 -> 57.2 ms, 2 -> 56.3 ms, compiler's choice (4) -> 56.7 ms
#endif
#if defined(__CUDA_ARCH__)
#define FLASH_SKIN_LOOP FLASH_PRAGMA(unroll FLASH_SKIN_UNROLL)
#else
#define FLASH_SKIN_LOOP
#endif
// approximate 1/sqrt(x) for normal x > 0: one MUFU.RSQ (rel. error 2^-22), no denormal fix-up code
FLASH_HD float rsqrt_fast(float x) {
#ifdef __CUDA_ARCH__
  float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
#else
  return 1.0f / sqrtf(x);
#endif
}

// ------------------------------------------------------------------ small algebra
FLASH_HD F3 f3(float x, float y, float z) { F3 r; r.x = x; r.y = y; r.z = z; return r; }
FLASH_HD F3 operator-(F3 a, F3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
FLASH_HD F3 operator+(F3 a, F3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
FLASH_HD F3 operator*(F3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
FLASH_HD float dot(F3 a, F3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
FLASH_HD F3 cross(F3 a, F3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
FLASH_HD F3 xyz(F4 a) { return f3(a.x, a.y, a.z); }
FLASH_HD D3 d3(double x, double y, double z) { D3 r; r.x = x; r.y = y; r.z = z; return r; }
FLASH_HD D3 operator-(D3 a, D3 b) { return d3(a.x - b.x, a.y - b.y, a.z - b.z); }
FLASH_HD D3 operator+(D3 a, D3 b) { return d3(a.x + b.x, a.y + b.y, a.z + b.z); }
FLASH_HD D3 operator*(D3 a, double s) { return d3(a.x * s, a.y * s, a.z * s); }
FLASH_HD double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
FLASH_HD D3 cross(D3 a, D3 b) { return d3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
FLASH_HD D3 xf_rot(const double* X, D3 p) { return d3(X[0] * p.x + X[1] * p.y + X[2] * p.z, X[3] * p.x + X[4] * p.y + X[5] * p.z, X[6] * p.x + X[7] * p.y + X[8] * p.z); }
FLASH_HD D3 xf_rotT(const double* X, D3 p) { return d3(X[0] * p.x + X[3] * p.y + X[6] * p.z, X[1] * p.x + X[4] * p.y + X[7] * p.z, X[2] * p.x + X[5] * p.y + X[8] * p.z); }
FLASH_HD D3 xf_t(const double* X) { return d3(X[9], X[10], X[11]); }
FLASH_HD D3 xf_apply(const double* X, D3 p) { return xf_rot(X, p) + xf_t(X); }
// XfD stored flat as 12 doubles: R[0..8], t[9..11]
FLASH_HD void xf_compose(const double* A, const double* B, double* C) {  // C = A*B (C may not alias)
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  D3 t = xf_apply(A, d3(B[9], B[10], B[11]));
  C[9] = t.x; C[10] = t.y; C[11] = t.z;
}
FLASH_HD void xf_from_rowmajor34(const double* m, double* X) {
  for (int i = 0; i < 3; ++i) { X[3 * i] = m[4 * i]; X[3 * i + 1] = m[4 * i + 1]; X[3 * i + 2] = m[4 * i + 2]; X[9 + i] = m[4 * i + 3]; }
}

// ------------------------------------------------------------------ kinematics (row A3/A4/A10)
// Joint transform of body b under the normalised configuration.  Revolute(axis) and
// QuaternionFloating q=[w x y z tx ty tz] as in src/models.jl:26,79,106.
FLASH_HD void joint_xf(const BodyC& B, const double* qn, double* J) {
  for (int i = 0; i < 12; ++i) J[i] = 0;
  J[0] = J[4] = J[8] = 1;
  if (B.jtype == JOINT_REVOLUTE) {
    double ax = B.axis[0], ay = B.axis[1], az = B.axis[2];
    double s, c;
#ifdef __CUDA_ARCH__
    sincos(qn[B.qoff], &s, &c);
#else
    s = sin(qn[B.qoff]); c = cos(qn[B.qoff]);
#endif
    double C = 1.0 - c;
    J[0] = c + C * ax * ax;      J[1] = C * ax * ay - s * az; J[2] = C * ax * az + s * ay;
    J[3] = C * ay * ax + s * az; J[4] = c + C * ay * ay;      J[5] = C * ay * az - s * ax;
    J[6] = C * az * ax - s * ay; J[7] = C * az * ay + s * ax; J[8] = c + C * az * az;
  } else if (B.jtype == JOINT_QUAT) {
    const double w = qn[B.qoff], x = qn[B.qoff + 1], y = qn[B.qoff + 2], z = qn[B.qoff + 3];
    J[0] = 1 - 2 * (y * y + z * z); J[1] = 2 * (x * y - w * z);     J[2] = 2 * (x * z + w * y);
    J[3] = 2 * (x * y + w * z);     J[4] = 1 - 2 * (x * x + z * z); J[5] = 2 * (y * z - w * x);
    J[6] = 2 * (x * z - w * y);     J[7] = 2 * (y * z + w * x);     J[8] = 1 - 2 * (x * x + y * y);
    J[9] = qn[B.qoff + 4]; J[10] = qn[B.qoff + 5]; J[11] = qn[B.qoff + 6];
  }
}

// Forward kinematics for one configuration.  x is the RAW state; quaternions are normalised here
// exactly where the reference does it (normalize!, src/gradientdescent.jl:19-30).  Writes into the
// per-config aux block: X_b (body->world), P_b (joint frameBefore->world), |quat_b|.
FLASH_HD void fk_bodies(const BodyC* __restrict__ body, int nb, int nq, const double* x, double* aux) {
  double qn[7 * MAX_BODIES > 64 ? 64 : 7 * MAX_BODIES];  // nq <= 64 enforced at model creation
  for (int k = 0; k < nq; ++k) qn[k] = x[k];
  for (int b = 0; b < nb; ++b) {
    const BodyC& B = body[b];
    double nrm = 1.0;
    if (B.jtype == JOINT_QUAT) {
      const double* q = x + B.qoff;
      nrm = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
      for (int k = 0; k < 4; ++k) qn[B.qoff + k] = q[k] / nrm;
    }
    aux[24 * nb + b] = nrm;
  }
  for (int b = 0; b < nb; ++b) {
    const BodyC& B = body[b];
    double j2p[12], J[12];
    xf_from_rowmajor34(B.j2p, j2p);
    double* P = aux + 12 * nb + 12 * b;
    if (B.parent >= 0) xf_compose(aux + 12 * B.parent, j2p, P);
    else for (int i = 0; i < 12; ++i) P[i] = j2p[i];
    joint_xf(B, qn, J);
    xf_compose(P, J, aux + 12 * b);
  }
}
FLASH_HD void fk_config(const ModelC& M, const double* x, double* aux) { fk_bodies(M.body, M.n_bodies, M.nq, x, aux); }

// ------------------------------------------------------------------ LU (RBF system, row A7)
// In-place LU with partial pivoting, row-major M x M.  `lane`/`nl` let a warp share the row
// updates (nl = 1 on the host); WSYNC must be a warp barrier on the device.
#ifdef __CUDA_ARCH__
#define FLASH_WSYNC() __syncwarp()
#else
#define FLASH_WSYNC() ((void)0)
#endif
FLASH_HD int lu_factor(double* A, double* piv, int M, int lane, int nl) {
  int ok = 1;
  for (int k = 0; k < M; ++k) {
    int p = k; double best = fabs(A[k * M + k]);
    for (int i = k + 1; i < M; ++i) { double v = fabs(A[i * M + k]); if (v > best) { best = v; p = i; } }
    if (best == 0.0) ok = 0;
    FLASH_WSYNC();
    if (p != k) for (int j = lane; j < M; j += nl) { double t = A[k * M + j]; A[k * M + j] = A[p * M + j]; A[p * M + j] = t; }
    if (lane == 0) piv[k] = (double)p;
    FLASH_WSYNC();
    const double inv = 1.0 / A[k * M + k];
    for (int i = k + 1 + lane; i < M; i += nl) {
      const double f = A[i * M + k] * inv;
      A[i * M + k] = f;
      for (int j = k + 1; j < M; ++j) A[i * M + j] -= f * A[k * M + j];
    }
    FLASH_WSYNC();
  }
  return ok;
}
FLASH_HD void lu_solve(const double* LU, const double* piv, double* b, int M) {
  for (int k = 0; k < M; ++k) { int p = (int)piv[k]; if (p != k) { double t = b[k]; b[k] = b[p]; b[p] = t; } }
  for (int i = 1; i < M; ++i) { double s = b[i]; for (int j = 0; j < i; ++j) s -= LU[i * M + j] * b[j]; b[i] = s; }
  for (int i = M - 1; i >= 0; --i) { double s = b[i]; for (int j = i + 1; j < M; ++j) s -= LU[i * M + j] * b[j]; b[i] = s / LU[i * M + i]; }
}

// Per-config preparation after fk_config: convex world->geometry transforms and, per skin, the
// world control points (src/Flash.jl:152-205), the Turk-O'Brien system (InterpolatingSurface,
// src/Flash.jl:212) factored in fp64, and its solution.  tab/aux are this config's blocks.
FLASH_HD void prepare_convex_one(const ModelC& M, int j, const double* aux, double* tab) {
  {
    const ConvexC& C = M.cvx[j];
    double S[12];
    if (C.body >= 0) xf_compose(aux + aux_X(M, C.body), C.frame, S);
    else for (int i = 0; i < 12; ++i) S[i] = C.frame[i];
    double* T = tab + 12 * j;   // y = A x + b,  A = R^T, b = -R^T t
    T[0] = S[0]; T[1] = S[3]; T[2] = S[6];
    T[3] = S[1]; T[4] = S[4]; T[5] = S[7];
    T[6] = S[2]; T[7] = S[5]; T[8] = S[8];
    D3 b = xf_rotT(S, d3(S[9], S[10], S[11]));
    T[9] = -b.x; T[10] = -b.y; T[11] = -b.z;
  }
}
FLASH_HD void prepare_convex(const ModelC& M, const double* aux, double* tab) {
  for (int j = 0; j < M.n_convex; ++j) prepare_convex_one(M, j, aux, tab);
}
// world position of control point i of skin K (src/Flash.jl:152-205)
FLASH_HD D3 skin_centre_one(const ModelC& M, const SkinC& K, int i, const double* x, const double* aux) {
  D3 p = d3(M.skin_pt[K.pt_off + i][0], M.skin_pt[K.pt_off + i][1], M.skin_pt[K.pt_off + i][2]);
  if (K.kind == SURF_DEFORMABLE_SKIN && i < K.nsurf) {  // point + deformation in the body frame (src/Flash.jl:176-181)
    const double* dd = x + M.nq + K.deform_off + 3 * i;
    p = p + d3(dd[0], dd[1], dd[2]);
  }
  const int b = M.skin_pt_body[K.pt_off + i];
  if (b >= 0) p = xf_apply(aux + aux_X(M, b), p);
  return p;
}
FLASH_HD void skin_centres(const ModelC& M, const SkinC& K, const double* x, const double* aux, double* tab) {
  double* c = tab + K.tab_off;
  for (int i = 0; i < K.npts; ++i) {
    const D3 p = skin_centre_one(M, K, i, x, aux);
    c[3 * i] = p.x; c[3 * i + 1] = p.y; c[3 * i + 2] = p.z;
  }
}
FLASH_HD void skin_assemble(const ModelC& M, const SkinC& K, const double* tab, double* A, int lane, int nl) {
  const double* c = tab + K.tab_off;
  const int n = K.npts, Mm = K.M;
  for (int e = lane; e < Mm * Mm; e += nl) {
    int i = e / Mm, j = e % Mm;
    double v = 0;
    if (i < n && j < n) {
      double dx = c[3 * i] - c[3 * j], dy = c[3 * i + 1] - c[3 * j + 1], dz = c[3 * i + 2] - c[3 * j + 2];
      double r2 = dx * dx + dy * dy + dz * dz;
      v = r2 * sqrt(r2);
    } else if (i < n) v = (j == n) ? 1.0 : c[3 * i + (j - n - 1)];
    else if (j < n) v = (i == n) ? 1.0 : c[3 * j + (i - n - 1)];
    A[e] = v;
  }
}
// rhs zeros(surface), -1 (skeleton) (src/Flash.jl:211); solution -> tab (w, a0, a)
FLASH_HD void skin_solve(const ModelC& M, const SkinC& K, const double* LU, const double* piv, double* tab) {
  double b[MAX_SKIN_M];
  const int n = K.npts;
  for (int i = 0; i < K.M; ++i) b[i] = (i < n && i >= K.nsurf) ? -1.0 : 0.0;
  lu_solve(LU, piv, b, K.M);
  double* w = tab + K.tab_off + 3 * n;
  for (int i = 0; i < n; ++i) w[i] = b[i];
  w[n] = b[n];
  for (int k = 0; k < 3; ++k) w[n + 1 + k] = (K.M == n + 4) ? b[n + 1 + k] : 0.0;
}

// ------------------------------------------------------------------ convex signed distance (row A8)
// Closest point to the ORIGIN on segment / triangle given in coordinates relative to the query
// point (Ericson, Real-Time Collision Detection 5.1.2 / 5.1.5), with barycentric weights whose
// exact zeros identify the supporting sub-simplex.
FLASH_HD F3 seg_closest0(F3 a, F3 b, float& wa, float& wb) {
  F3 ab = b - a;
  float den = dot(ab, ab);
  float t = den > 0.f ? -dot(a, ab) / den : 0.f;
  if (t <= 0.f) { wa = 1.f; wb = 0.f; return a; }
  if (t >= 1.f) { wa = 0.f; wb = 1.f; return b; }
  wa = 1.f - t; wb = t;
  return a + ab * t;
}
// cold path of tri_closest0 (collinear triangle: best of the three edges), kept out of line so that the five
// inlined copies of the triangle solver do not each carry three segment solvers (instruction-cache footprint)
#if defined(FLASH_COLD_NOINLINE)
FLASH_HDN
#else
FLASH_HD
#endif
F3 tri_degenerate0(F3 a, F3 b, F3 c, float& wa, float& wb, float& wc) {
  float u0, u1, v0, v1, w0, w1;
  F3 q0 = seg_closest0(a, b, u0, u1), q1 = seg_closest0(a, c, v0, v1), q2 = seg_closest0(b, c, w0, w1);
  float e0 = dot(q0, q0), e1 = dot(q1, q1), e2 = dot(q2, q2);
  if (e0 <= e1 && e0 <= e2) { wa = u0; wb = u1; wc = 0.f; return q0; }
  if (e1 <= e2) { wa = v0; wb = 0.f; wc = v1; return q1; }
  wa = 0.f; wb = w0; wc = w1; return q2;
}
FLASH_HD F3 tri_closest0(F3 a, F3 b, F3 c, float& wa, float& wb, float& wc) {
  F3 ab = b - a, ac = c - a;
  float d1 = -dot(ab, a), d2 = -dot(ac, a);
  if (d1 <= 0.f && d2 <= 0.f) { wa = 1.f; wb = 0.f; wc = 0.f; return a; }
  float d3 = -dot(ab, b), d4 = -dot(ac, b);
  if (d3 >= 0.f && d4 <= d3) { wa = 0.f; wb = 1.f; wc = 0.f; return b; }
  float vc = d1 * d4 - d3 * d2;
  if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { float v = d1 / (d1 - d3); wa = 1.f - v; wb = v; wc = 0.f; return a + ab * v; }
  float d5 = -dot(ab, c), d6 = -dot(ac, c);
  if (d6 >= 0.f && d5 <= d6) { wa = 0.f; wb = 0.f; wc = 1.f; return c; }
  float vb = d5 * d2 - d1 * d6;
  if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { float w = d2 / (d2 - d6); wa = 1.f - w; wb = 0.f; wc = w; return a + ac * w; }
  float va = d3 * d6 - d5 * d4;
  if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) { float w = (d4 - d3) / ((d4 - d3) + (d5 - d6)); wa = 0.f; wb = 1.f - w; wc = w; return b + (c - b) * w; }
  float den = va + vb + vc;
  if (!(den > 0.f)) return tri_degenerate0(a, b, c, wa, wb, wc);   // degenerate (collinear) triangle: best of the three edges
  float inv = 1.f / den;
  float v = vb * inv, w = vc * inv;
  wa = 1.f - v - w; wb = v; wc = w;
  return a + ab * v + ac * w;
}

// GJK working simplex: up to three hull vertices (coordinates relative to the query point).
struct Simplex { F3 a, b, c; int ia, ib, ic, ns; };
FLASH_HD void simplex_push(Simplex& S, F3 p, int j) {
  if (S.ns == 0) { S.a = p; S.ia = j; } else if (S.ns == 1) { S.b = p; S.ib = j; } else { S.c = p; S.ic = j; }
  ++S.ns;
}
// keep the vertices with non-zero barycentric weight (at least one)
FLASH_HD void simplex_from(Simplex& S, F3 p0, int j0, float u0, F3 p1, int j1, float u1, F3 p2, int j2, float u2) {
  S.ns = 0;
  if (u0 > 0.f) simplex_push(S, p0, j0);
  if (u1 > 0.f) simplex_push(S, p1, j1);
  if (u2 > 0.f) simplex_push(S, p2, j2);
  if (S.ns == 0) simplex_push(S, p2, j2);
}

// ---- support direction map (per hull): a cube map of DIRMAP_R x DIRMAP_R cells per face over directions d.
//   voff/vlist  (u16 offsets / u8 lists)  EVERY vertex that is the support vertex for SOME direction of
//                      the cell (host: normal-cone / cell clipping with margins) -- the support query over
//                      the list (2-3 vertices on the IRB140 hulls) is exact and replaces the 56-vertex scan.
// (A face table of the same shape that replaced the 100-plane scan by one plane test was measured and
//  dropped: the most-violated plane is a much better GJK start -- 1.0 vs 1.3 iterations per point, and the
//  iterations, not the scan, are what costs)
constexpr int DIRMAP_R = 8;
constexpr int DIRMAP_CELLS = 6 * DIRMAP_R * DIRMAP_R;
FLASH_HD int dir_cell(F3 d) {
  // branch-free: major axis by selects, one (approximate on the device: 2 ulp, far inside the margin the
  // host grows every cell by) reciprocal, clamps in floating point
  const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
  const bool X = ax >= ay && ax >= az, Y = !X && ay >= az;
  const float m = X ? d.x : (Y ? d.y : d.z), u = X ? d.y : (Y ? d.z : d.x), v = X ? d.z : (Y ? d.x : d.y);
  const int face = (X ? 0 : (Y ? 2 : 4)) + (m < 0.f ? 1 : 0);
#ifdef __CUDA_ARCH__
  const float inv = __fdividef(0.5f * DIRMAP_R, fmaxf(fabsf(m), 1e-30f));
#else
  const float inv = (0.5f * DIRMAP_R) / fmaxf(fabsf(m), 1e-30f);
#endif
  // (u/|m| + 1) R/2 in [0, R] -> cell
  const float fu = fminf(fmaxf(fmaf(u, inv, 0.5f * DIRMAP_R), 0.f), DIRMAP_R - 0.5f);
  const float fv = fminf(fmaxf(fmaf(v, inv, 0.5f * DIRMAP_R), 0.f), DIRMAP_R - 0.5f);
  return (face * DIRMAP_R + (int)fv) * DIRMAP_R + (int)fu;
}

// ---- geometry blob layout (per hull), chosen so that the scans run on packed fp32 FMAs (FFMA2, sm_100):
//   planes, pair k = planes 2k, 2k+1:  F4 {nx0, nx1, ny0, ny1}, F4 {nz0, nz1, -o0, -o1}
//   vertices, pair k:                  F4 {x0, x1, y0, y1} (VA) and float2 {z0, z1} (VZ, a flat float array)
// counts are padded to a multiple of 8; padding planes can never be the maximum, padding vertices are
// copies of vertex 0 (never a strictly better support).
struct HullView { const F4* planes; const uint32_t* ptri; const F4* va; const float* vz; int np, nv;
                  const unsigned short* voff; const unsigned char* vlist; const unsigned short* pgrp;
                  const F4* bn; const uint32_t* bk; const F4* bp; const F4* bv; const F4* be; int bsp_on; };
FLASH_HD F3 plane_normal(const F4* P, int f) { const float* a = reinterpret_cast<const float*>(P + (f & ~1)); const int h = f & 1; return f3(a[h], a[2 + h], a[4 + h]); }
FLASH_HD F3 vert_at(const HullView& V, int j) { const float* a = reinterpret_cast<const float*>(V.va + (j >> 1)); const int h = j & 1; return f3(a[h], a[2 + h], V.vz[j]); }

// eight plane distances n.y - o / eight support values q.v; the device versions issue one FFMA2 per
// two lanes of work, the host versions are the same fused multiply-adds one at a time (bit-identical).
#ifdef __CUDA_ARCH__
__device__ __forceinline__ unsigned long long pk2(float lo, float hi) { unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(unsigned long long v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
  unsigned long long r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
#endif
FLASH_HD void planes8(const F4* __restrict__ P, F3 y, float sv[8]) {
#if defined(__CUDA_ARCH__) && !defined(FLASH_NO_FFMA2)
  const unsigned long long yx = pk2(y.x, y.x), yy = pk2(y.y, y.y), yz = pk2(y.z, y.z);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const F4 a = P[2 * k], b = P[2 * k + 1];
    unsigned long long r = fma2(pk2(b.x, b.y), yz, pk2(b.z, b.w));
    r = fma2(pk2(a.z, a.w), yy, r);
    r = fma2(pk2(a.x, a.y), yx, r);
    upk2(r, sv[2 * k], sv[2 * k + 1]);
  }
#else
  for (int k = 0; k < 4; ++k) {
    const F4 a = P[2 * k], b = P[2 * k + 1];
    sv[2 * k] = fmaf(a.x, y.x, fmaf(a.z, y.y, fmaf(b.x, y.z, b.z)));
    sv[2 * k + 1] = fmaf(a.y, y.x, fmaf(a.w, y.y, fmaf(b.y, y.z, b.w)));
  }
#endif
}
FLASH_HD void verts8(const F4* __restrict__ VA, const float* __restrict__ VZ, F3 q, float sv[8]) {
#if defined(__CUDA_ARCH__) && !defined(FLASH_NO_FFMA2)
  const unsigned long long qx = pk2(q.x, q.x), qy = pk2(q.y, q.y), qz = pk2(q.z, q.z);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const F4 a = VA[k];
    const float2 z = *reinterpret_cast<const float2*>(VZ + 2 * k);
    unsigned long long r = mul2(pk2(z.x, z.y), qz);
    r = fma2(pk2(a.z, a.w), qy, r);
    r = fma2(pk2(a.x, a.y), qx, r);
    upk2(r, sv[2 * k], sv[2 * k + 1]);
  }
#else
  for (int k = 0; k < 4; ++k) {
    const F4 a = VA[k];
    sv[2 * k] = fmaf(q.x, a.x, fmaf(q.y, a.z, q.z * VZ[2 * k]));
    sv[2 * k + 1] = fmaf(q.x, a.y, fmaf(q.y, a.w, q.z * VZ[2 * k + 1]));
  }
#endif
}

// Signed distance from y (geometry frame) to the convex hull given by its supporting planes and
// vertices.  Inside (all plane distances <= 0): the largest plane distance, normal = that plane's.
// Outside: exact Euclidean distance by GJK on the vertex set, started from a triangle of the
// most-violated face (EnhancedGJK `gjk!` is the reference algorithm, src/Flash.jl:238-243; its
// inside value is iteration-path dependent and replaced by the exact depth -- DESIGN.md).
// nrm receives grad_y(distance).
//
// The function is split into the two scans (plane maximum, support vertex) and the per-point pieces
// between them, because the main kernel runs the scans warp-cooperatively (lanes = planes / vertices,
// flash_kernels.cuh hull_sd_warp) around the SAME per-point pieces; this serial composition is what
// the ray marcher, skin_eval, grid baking and the host replay use.

// [region: plane scan] largest plane distance D and the lowest plane index attaining it.  Returns false
// (and leaves fs undefined) as soon as D exceeds `cutoff`: D is a lower bound of the distance.
FLASH_HD bool hull_plane_scan(const HullView& V, F3 y, float cutoff, float& D_out, int& fs_out) {
  const F4* __restrict__ planes = V.planes;
  float D = -3.0e38f; int fg = 0;
  int f = 0;
  for (; f < V.np; f += 8) {
    float sv[8];
    planes8(planes + f, y, sv);
    const float m = fmaxf(fmaxf(fmaxf(sv[0], sv[1]), fmaxf(sv[2], sv[3])), fmaxf(fmaxf(sv[4], sv[5]), fmaxf(sv[6], sv[7])));
    fg = m > D ? f : fg;
    D = fmaxf(D, m);
    if (D > cutoff) { D_out = D; return false; }
  }
  float sv[8];                // index inside the winning group: recomputed once instead of tracked per group
  planes8(planes + fg, y, sv);
  int k = 7;
#pragma unroll
  for (int t = 6; t >= 0; --t) k = (sv[t] == D) ? t : k;
  D_out = D; fs_out = fg + k;
  return true;
}
// [region: plane hint] The host clusters the planes into groups of 8 by normal direction and records,
// per direction cell of y (geometry origin = hull centroid), the two groups that hold the most-violated
// plane for (nearly) every OUTSIDE point of that cell (flash_host.h build_plane_hints: upper envelopes of
// the plane distances along sampled rays).  The map is a HINT, never trusted for the result:
//   * any plane with n.y - o > 0 proves y outside, and its distance is a lower bound of the true one
//     (valid for the cutoff test and as the "certified face" shortcut of hull_gjk_start);
//   * the distance itself comes from GJK, which terminates on its own exact criterion whatever face it
//     started from (a missed most-violated plane costs an extra iteration, not accuracy);
//   * if no hinted plane is violated the point is inside or the hint failed: the full scan decides.
// Returns 0: D > cutoff (cannot be the minimum), 1: outside with (D, fs), 2: undecided.
FLASH_HD float max8(const float sv[8]) { return fmaxf(fmaxf(fmaxf(sv[0], sv[1]), fmaxf(sv[2], sv[3])), fmaxf(fmaxf(sv[4], sv[5]), fmaxf(sv[6], sv[7]))); }
FLASH_HD int hull_plane_hint(const HullView& V, F3 y, float cutoff, float& D_out, int& fs_out) {
  const unsigned gg = V.pgrp[dir_cell(y)];
  const int g0 = (int)(gg & 255u) * 8, g1 = (int)(gg >> 8) * 8;
  float sv[8];
  planes8(V.planes + g0, y, sv);
  float D = max8(sv);
  int gw = g0;
  if (D > cutoff) { D_out = D; return 0; }        // most rejections need the first group only
  if (g1 != g0) {
    planes8(V.planes + g1, y, sv);
    const float m = max8(sv);
    gw = m > D ? g1 : g0;
    D = fmaxf(D, m);
    if (D > cutoff) { D_out = D; return 0; }
  }
  D_out = D;
  if (!(D > 0.f)) return 2;
  planes8(V.planes + gw, y, sv);                   // index inside the winning group: recomputed for the accepted points only
  int k = 7;
#pragma unroll
  for (int t = 6; t >= 0; --t) k = (sv[t] == D) ? t : k;
  fs_out = gw + k;
  return 1;
}
// [region: support scan] lowest vertex index minimising q.v
FLASH_HD int hull_support_scan(const HullView& V, F3 q) {
  float best = 3.0e38f; int vg = 0;
  for (int v = 0; v < V.nv; v += 8) {   // groups of 8 (padded with copies of vertex 0)
    float sv[8];
    verts8(V.va + (v >> 1), V.vz + v, q, sv);
    const float m = fminf(fminf(fminf(sv[0], sv[1]), fminf(sv[2], sv[3])), fminf(fminf(sv[4], sv[5]), fminf(sv[6], sv[7])));
    vg = m < best ? v : vg;
    best = fminf(best, m);
  }
  float sv[8];
  verts8(V.va + (vg >> 1), V.vz + vg, q, sv);
  int k = 7;
#pragma unroll
  for (int t = 6; t >= 0; --t) k = (sv[t] == best) ? t : k;
  return vg + k;
}
// [region: support map] exact support vertex (lowest index minimising q.v) from the direction map: only
// the vertices whose normal cone meets the cell of -q can be the support (host: build_dirmaps)
FLASH_HD int hull_support_map(const HullView& V, F3 q) {
  const int cell = dir_cell(f3(-q.x, -q.y, -q.z));
  const int b = V.voff[cell], e = V.voff[cell + 1];
  float best = 3.0e38f; int iw = 0;
  for (int t = b; t < e; ++t) {       // lists are in increasing vertex order: strict < keeps the lowest index on ties
    const int j = V.vlist[t];
    const F3 p = vert_at(V, j);
    const float sv = fmaf(q.x, p.x, fmaf(q.y, p.y, q.z * p.z));
    if (sv < best) { best = sv; iw = j; }
  }
  return iw;
}
// [region: start triangle] closest point of face fs's triangle; true = it is interior to the triangle,
// i.e. the orthogonal projection of y onto the supporting plane, which lies on the hull: it IS the
// closest point (no tolerance involved) and no support scan is needed to certify it.
FLASH_HD bool hull_gjk_start(const HullView& V, F3 y, int fs, Simplex& S, F3& q) {
  const uint32_t t = V.ptri[fs];
  const int j0 = t & 255, j1 = (t >> 8) & 255, j2 = (t >> 16) & 255;
  F3 p0 = vert_at(V, j0) - y, p1 = vert_at(V, j1) - y, p2 = vert_at(V, j2) - y;
  float u0, u1, u2;
  q = tri_closest0(p0, p1, p2, u0, u1, u2);
  simplex_from(S, p0, j0, u0, p1, j1, u1, p2, j2, u2);
  return S.ns == 3;
}
// [region: simplex update] one GJK step after the support scan returned vertex iw for direction q.
// Returns false when q is already the closest point (iw supports the simplex / no descent left).
FLASH_HD bool hull_gjk_update(const HullView& V, F3 y, int iw, Simplex& S, F3& q) {
  if (iw == S.ia || (S.ns > 1 && iw == S.ib) || (S.ns > 2 && iw == S.ic)) return false;  // already supporting
  const float dd = dot(q, q);
  F3 w = vert_at(V, iw) - y;
  if (dd - dot(q, w) <= 1e-7f * dd) return false;                                          // no descent left
  float u0, u1, u2;
  if (S.ns == 1) {
    q = seg_closest0(S.a, w, u0, u1);
    simplex_from(S, S.a, S.ia, u0, w, iw, u1, w, iw, 0.f);
  } else if (S.ns == 2) {
    q = tri_closest0(S.a, S.b, w, u0, u1, u2);
    simplex_from(S, S.a, S.ia, u0, S.b, S.ib, u1, w, iw, u2);
  } else {
    // tetrahedron (a,b,c,w): y is outside the hull, so the closest point lies on one of the
    // three faces that contain the new vertex.  (A rolled loop over the faces -- one copy of the triangle
    // solver instead of four, 5088 instead of 6072 instructions -- was measured 10% SLOWER: the dynamic
    // indexing of the simplex spills.)
    float x0, x1, x2, y0, y1, y2, z0, z1, z2;
    F3 q0 = tri_closest0(S.a, S.b, w, x0, x1, x2);
    F3 q1 = tri_closest0(S.a, S.c, w, y0, y1, y2);
    F3 q2 = tri_closest0(S.b, S.c, w, z0, z1, z2);
    float e0 = dot(q0, q0), e1 = dot(q1, q1), e2 = dot(q2, q2);
    Simplex T = S;
    if (e0 <= e1 && e0 <= e2) { q = q0; simplex_from(S, T.a, T.ia, x0, T.b, T.ib, x1, w, iw, x2); }
    else if (e1 <= e2) { q = q1; simplex_from(S, T.a, T.ia, y0, T.c, T.ic, y1, w, iw, y2); }
    else { q = q2; simplex_from(S, T.b, T.ib, z0, T.c, T.ic, z1, w, iw, z2); }
  }
  return true;
}
// distance and its gradient from the converged q (closest point - y)
FLASH_HD float hull_gjk_finish(const HullView& V, F3 q, float D, int fs, F3& nrm) {
  const float dd = dot(q, q);
  if (!(dd > 0.f)) { nrm = plane_normal(V.planes, fs); return D; }   // numerically on the surface
#ifdef __CUDA_ARCH__
  float inv = rsqrtf(dd);
  inv = inv * (1.5f - 0.5f * dd * inv * inv);   // one Newton step: ~1 ulp
#else
  float inv = 1.0f / sqrtf(dd);
#endif
  nrm = q * (-inv);
  return dd * inv;
}

// kind_out (may be null): 0 = rejected by the cutoff, 1 = inside (fs = the plane of largest distance), 2 = outside with
// the closest point interior to the start triangle of plane fs, 3 = outside with GJK's terminal simplex in S_out.
FLASH_HD float hull_sd_ex(const HullView& V, F3 y, float cutoff, F3& nrm, int* kind_out, int* fs_out, Simplex* S_out) {
  float D = 0.f; int fs = 0;
  if (kind_out) *kind_out = 0;
#ifndef FLASH_NO_PLANEHINT
  const int hint = hull_plane_hint(V, y, cutoff, D, fs);
  FLASH_COUNT(plane_tests, 16);
  if (hint == 0) { nrm = f3(0.f, 0.f, 0.f); return 3.0e38f; }
  if (hint == 2) {
    FLASH_COUNT(plane_tests, V.np);
    if (!hull_plane_scan(V, y, cutoff, D, fs)) { nrm = f3(0.f, 0.f, 0.f); return 3.0e38f; }
  }
#else
  if (!hull_plane_scan(V, y, cutoff, D, fs)) { nrm = f3(0.f, 0.f, 0.f); return 3.0e38f; }   // cannot be the minimum: never a candidate value
  FLASH_COUNT(plane_tests, V.np);
#endif
  FLASH_COUNT(narrow, 1);
  if (fs_out) *fs_out = fs;
  if (D <= 0.f) { FLASH_COUNT(inside, 1); if (kind_out) *kind_out = 1; nrm = plane_normal(V.planes, fs); return D; }
  Simplex S;
  F3 q;
  const bool certified = hull_gjk_start(V, y, fs, S, q);
  for (int it = 0; it < GJK_MAX_ITER && !certified; ++it) {
    if (!(dot(q, q) > 0.f)) break;
    // support vertex towards the query point: minimise q.(v - y)  <=>  minimise q.v
    FLASH_COUNT(gjk_iters, 1);
#ifndef FLASH_NO_DIRMAP
    const int iw = hull_support_map(V, q);
#else
    FLASH_COUNT(vert_tests, V.nv);
    const int iw = hull_support_scan(V, q);
#endif
    if (!hull_gjk_update(V, y, iw, S, q)) break;
  }
  if (kind_out) *kind_out = certified ? 2 : 3;
  if (S_out) *S_out = S;
  return hull_gjk_finish(V, q, D, fs, nrm);
}
FLASH_HD float hull_sd(const HullView& V, F3 y, float cutoff, F3& nrm) { return hull_sd_ex(V, y, cutoff, nrm, nullptr, nullptr, nullptr); }

// ---- closest-feature BSP (flash_bsp.h): one plane test per level locates the cell of y, whose feature gives the
// signed distance and its gradient in closed form.  Leaf word: 0x8000 | kind << 13 | index, kind 0 = face (plane
// index), 1 = vertex, 2 = edge (edge table: unit direction + first vertex).  Lane-uniform: every lane runs the same
// ~9 instructions per level, whatever its path.
FLASH_HD unsigned bsp_locate(const F4* __restrict__ nodes, const uint32_t* __restrict__ kids, F3 y) {
  unsigned k = 0u;
  do {
#ifdef __CUDA_ARCH__
    const float4 pp = __ldg(reinterpret_cast<const float4*>(nodes) + k);
    F4 p; p.x = pp.x; p.y = pp.y; p.z = pp.z; p.w = pp.w;
    const uint32_t w = __ldg(kids + k);
#else
    const F4 p = nodes[k];
    const uint32_t w = kids[k];
#endif
    const float s = fmaf(p.x, y.x, fmaf(p.y, y.y, fmaf(p.z, y.z, p.w)));
    k = s > 0.f ? (w >> 16) : (w & 0xffffu);
  } while (!(k & 0x8000u));
  return k;
}
FLASH_HD float bsp_feature_sd(const HullView& V, unsigned leaf, F3 y, F3& nrm) {
  const unsigned kind = (leaf >> 13) & 3u, idx = leaf & 0x1fffu;
  if (kind == 0u) {
    const F4 p = V.bp[idx];
    nrm = f3(p.x, p.y, p.z);
    return fmaf(p.x, y.x, fmaf(p.y, y.y, fmaf(p.z, y.z, p.w)));
  }
  F3 u = f3(0.f, 0.f, 0.f);
  unsigned ia = idx;
  if (kind == 2u) {
    const F4 e = V.be[idx];
    u = f3(e.x, e.y, e.z);
#ifdef __CUDA_ARCH__
    ia = __float_as_uint(e.w);
#else
    memcpy(&ia, &e.w, 4);
#endif
  }
  const F4 a = V.bv[ia];
  F3 r = y - f3(a.x, a.y, a.z);
  const float t = dot(r, u);
  r = f3(fmaf(-t, u.x, r.x), fmaf(-t, u.y, r.y), fmaf(-t, u.z, r.z));   // vertex: u = 0
  const float dd = dot(r, r);
  if (!(dd > 0.f)) { nrm = f3(0.f, 0.f, 0.f); return 0.f; }                // on the feature itself
#ifdef __CUDA_ARCH__
  float inv = rsqrtf(dd);
  inv = inv * (1.5f - 0.5f * dd * inv * inv);   // one Newton step: ~1 ulp
#else
  const float inv = 1.0f / sqrtf(dd);
#endif
  nrm = r * inv;
  return dd * inv;
}
FLASH_HD float hull_sd_bsp(const HullView& V, F3 y, F3& nrm) { return bsp_feature_sd(V, bsp_locate(V.bn, V.bk, y), y, nrm); }

// ---- group feature certification ---------------------------------------------------------------------------------
// The 32 points one warp evaluates together are Morton neighbours inside a ball B(c, rho) (points_groups_kernel).
// If the whole ball lies in the Voronoi region of ONE feature of the hull -- a face, an edge or a vertex -- every
// point's distance is a closed form of that feature and the per-point narrow phase (plane hint, start triangle,
// GJK iterations, support queries) is not needed at all.  The cost kernel's group pre-pass (one thread per group)
// runs the narrow phase once at the ball centre, takes the closest feature from it and certifies it for the ball
// by the projection theorem -- x* in H is the closest point of x iff (x - x*).(z - x*) <= 0 for every z in H -- whose
// conditions are LINEAR in x for a fixed feature, so their maximum over a ball is the value at the centre plus rho
// times a norm:
//   vertex v        (x - v).(v_k - v) <= 0 for every hull vertex k
//   edge (a, b)     0 <= (x - a).e <= |b - a| (e the unit direction), and (x - a).w_k <= 0 with w_k the part of
//                   v_k - a orthogonal to e
//   face f          the projection x' of x onto the plane of f lies in H: n_g.x' - o_g <= 0 for every plane g (points
//                   outside), and f is the plane of largest distance: n_f.x - o_f >= n_g.x - o_g (points inside)
// A feature is a u32: type (bits 0-1: 0 none, 1 vertex, 2 edge, 3 face) | FEAT_CERT | ia << 8 | ib << 16 (vertices) or
// plane slot << 8 (face).  All tests carry an absolute slack for the fp32 centre transform; a point that is outside
// its certified region by a rounding error delta gets a distance that is wrong by O(delta^2 / d) only.
// Without FEAT_CERT the word is a HINT: the vertices (ia, ib, ic in bits 8-31, type = their number) of the simplex GJK
// ended on at the ball centre.  The points of the group start their own GJK from it (hull_sd_warm): no plane hint, no
// start triangle, and -- GJK's termination test being independent of the start -- the same exact distance.
enum { FEAT_NONE = 0, FEAT_VERTEX = 1, FEAT_EDGE = 2, FEAT_FACE = 3, FEAT_CERT = 4 };
FLASH_HD unsigned feat_hint(const Simplex& S) {
  return (unsigned)S.ns | ((unsigned)S.ia << 8) | ((unsigned)(S.ns > 1 ? S.ib : 0) << 16) | ((unsigned)(S.ns > 2 ? S.ic : 0) << 24);
}
FLASH_HD unsigned hull_ball_feature(const HullView& V, F3 yc, float rho, float& d_centre) {
  F3 n; int kind = 0, fs = 0; Simplex S;
  S.ns = 0; S.ia = S.ib = S.ic = 0; S.a = S.b = S.c = f3(0.f, 0.f, 0.f);
  const float D = hull_sd_ex(V, yc, 3.0e38f, n, &kind, &fs, &S);
  d_centre = D;
  const float r = rho * (1.f + 1e-5f) + 1e-5f;   // ball radius + slack for the fp32 transform of the centre
  // (a hint is only safe when the whole ball is outside the hull: GJK on the vertices cannot see penetration)
  const unsigned hint = (kind == 2 || kind == 3) && S.ns >= 1 && S.ns <= 3 && D - r > 0.f ? feat_hint(S) : (unsigned)FEAT_NONE;
  if (kind == 3 && S.ns == 3) {
    // GJK ended in the interior of a triangle of hull vertices: it lies on a supporting plane, i.e. on the face
    // whose plane is the farthest one from yc (every plane distance is a lower bound of the true distance)
    float Dp; int fp;
    if (!hull_plane_scan(V, yc, 3.0e38f, Dp, fp)) return hint;
    fs = fp; kind = 2;
  }
  if (kind == 1 || kind == 2) {
    // face fs
    const F3 nf = plane_normal(V.planes, fs);
    if (!(dot(nf, nf) > 0.5f)) return hint;
    const float* pf = reinterpret_cast<const float*>(V.planes + (fs & ~1));
    const float df = fmaf(nf.x, yc.x, fmaf(nf.y, yc.y, fmaf(nf.z, yc.z, pf[6 + (fs & 1)])));   // n_f.yc - o_f
    const bool dips_inside = df - r < 0.f, pokes_outside = df + r > 0.f;
    for (int g = 0; g < V.np; ++g) {
      if (g == fs) continue;
      const float* pg = reinterpret_cast<const float*>(V.planes + (g & ~1)); const int h = g & 1;
      const F3 ng = f3(pg[h], pg[2 + h], pg[4 + h]);
      const float dg = fmaf(ng.x, yc.x, fmaf(ng.y, yc.y, fmaf(ng.z, yc.z, pg[6 + h])));
      const float c = dot(ng, nf);
      // outside points: projection onto plane f stays inside plane g
      if (pokes_outside && !(dg - df * c + r * sqrtf(fmaxf(1.f - c * c, 0.f) + 1e-6f) <= 0.f)) return hint;
      // inside points: plane f has the largest distance
      if (dips_inside && !(df - dg - r * sqrtf(fmaxf(2.f - 2.f * c, 0.f) + 1e-6f) >= 0.f)) return hint;
    }
    return FEAT_FACE | FEAT_CERT | ((unsigned)fs << 8);
  }
  // Vertex and edge regions are NOT certified: with the warm start their points cost one closest-point update and
  // one support query, and the certificate (a pass over all vertices, or adjacency tables) costs the group kernel
  // more than the closed form saves (measured: K5 group_kernel 4.7 ms with the certificates, 24 % of the groups
  // certified).  Faces are: a patch of depth points lying ON a face straddles the surface, where GJK cannot start.
  return hint;
}
// GJK started from the group's hint simplex.  Returns false when the point turns out to be on / inside the hull
// (q = 0): the caller takes the regular path, whose plane scan gives the depth.
FLASH_HD bool hull_sd_warm(const HullView& V, unsigned feat, F3 y, float& d, F3& nrm) {
  const int ns = (int)(feat & 3u);
  const int ia = (int)((feat >> 8) & 255u), ib = (int)((feat >> 16) & 255u), ic = (int)(feat >> 24);
  Simplex S; F3 q;
  const F3 pa = vert_at(V, ia) - y;
  S.ns = 1; S.a = pa; S.ia = ia; S.b = pa; S.ib = ia; S.c = pa; S.ic = ia;
  q = pa;
  if (ns == 2) {
    const F3 pb = vert_at(V, ib) - y;
    float u0, u1;
    q = seg_closest0(pa, pb, u0, u1);
    simplex_from(S, pa, ia, u0, pb, ib, u1, pb, ib, 0.f);
  } else if (ns == 3) {
    const F3 pb = vert_at(V, ib) - y, pc = vert_at(V, ic) - y;
    float u0, u1, u2;
    q = tri_closest0(pa, pb, pc, u0, u1, u2);
    simplex_from(S, pa, ia, u0, pb, ib, u1, pc, ic, u2);
  }
  FLASH_COUNT(warm, 1);
  for (int it = 0; it < GJK_MAX_ITER; ++it) {
    if (!(dot(q, q) > 0.f)) return false;
    FLASH_COUNT(gjk_iters, 1);
    const int iw = hull_support_map(V, q);
    if (!hull_gjk_update(V, y, iw, S, q)) break;
  }
  const float dd = dot(q, q);
  if (!(dd > 0.f)) return false;
#ifdef __CUDA_ARCH__
  float inv = rsqrtf(dd);
  inv = inv * (1.5f - 0.5f * dd * inv * inv);
#else
  const float inv = 1.0f / sqrtf(dd);
#endif
  nrm = q * (-inv);
  d = dd * inv;
  return true;
}
// distance and gradient of y to a certified feature
FLASH_HD float hull_feature_eval(const HullView& V, unsigned feat, F3 y, F3& nrm) {
  const unsigned type = feat & 3u;
  if (type == FEAT_FACE) {
    const int fs = (int)(feat >> 8);
    const float* pf = reinterpret_cast<const float*>(V.planes + (fs & ~1)); const int h = fs & 1;
    nrm = f3(pf[h], pf[2 + h], pf[4 + h]);
    return fmaf(nrm.x, y.x, fmaf(nrm.y, y.y, fmaf(nrm.z, y.z, pf[6 + h])));
  }
  F3 u = y - vert_at(V, (int)((feat >> 8) & 255u));
  if (type == FEAT_EDGE) {
    const F3 e = vert_at(V, (int)((feat >> 16) & 255u)) - vert_at(V, (int)((feat >> 8) & 255u));
    u = u - e * (dot(u, e) / dot(e, e));
  }
  const float dd = dot(u, u);
#ifdef __CUDA_ARCH__
  float inv = rsqrtf(dd);
  inv = inv * (1.5f - 0.5f * dd * inv * inv);
#else
  const float inv = 1.0f / sqrtf(dd);
#endif
  nrm = u * inv;
  return dd * inv;
}

// conservative lower bound of the signed distance from the AABB of the hull
FLASH_HD float aabb_lower_bound(const float* lo, const float* hi, float depth_bound, F3 y) {
  float dx = fmaxf(fmaxf(lo[0] - y.x, y.x - hi[0]), 0.f);
  float dy = fmaxf(fmaxf(lo[1] - y.y, y.y - hi[1]), 0.f);
  float dz = fmaxf(fmaxf(lo[2] - y.z, y.z - hi[2]), 0.f);
  float d2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
  if (d2 > 0.f) return (d2 * rsqrt_fast(fmaxf(d2, 1e-30f))) * (1.f - 4e-6f) - 4e-6f;
  return -depth_bound;
}

// world point -> geometry frame, fp64 FMA chain rounded once to fp32 (DESIGN.md "precision")
FLASH_HD F3 to_geometry(const double* __restrict__ T, double x, double y, double z) {
  double a = fma(T[0], x, fma(T[1], y, fma(T[2], z, T[9])));
  double b = fma(T[3], x, fma(T[4], y, fma(T[5], z, T[10])));
  double c = fma(T[6], x, fma(T[7], y, fma(T[8], z, T[11])));
  return f3((float)a, (float)b, (float)c);
}

// fp32 version for the culling bound only (error < 2e-6 m for |x| < 4 m; the bound subtracts 4e-6)
FLASH_HD F3 to_geometry_f32(const float* __restrict__ T, float px, float py, float pz) {
  return f3(fmaf(T[0], px, fmaf(T[1], py, fmaf(T[2], pz, T[9]))), fmaf(T[3], px, fmaf(T[4], py, fmaf(T[5], pz, T[10]))),
            fmaf(T[6], px, fmaf(T[7], py, fmaf(T[8], pz, T[11]))));
}

// trilinear sample of a body-frame SDF grid with analytic gradient, manual fp32 (no texture unit)
FLASH_HD bool grid_sd(const ConvexC& C, const float* __restrict__ pool, F3 y, float& d, F3& nrm) {
  float gx = (y.x - C.g_org[0]) * C.g_inv_h, gy = (y.y - C.g_org[1]) * C.g_inv_h, gz = (y.z - C.g_org[2]) * C.g_inv_h;
  if (!(gx >= 0.f && gy >= 0.f && gz >= 0.f && gx < (float)(C.g_dim[0] - 1) && gy < (float)(C.g_dim[1] - 1) && gz < (float)(C.g_dim[2] - 1))) return false;
  int ix = (int)gx, iy = (int)gy, iz = (int)gz;
  float fx = gx - (float)ix, fy = gy - (float)iy, fz = gz - (float)iz;
  const float* g = pool + C.g_off + ((long long)iz * C.g_dim[1] + iy) * C.g_dim[0] + ix;
  const long long sy = C.g_dim[0], sz = (long long)C.g_dim[0] * C.g_dim[1];
  float c000 = g[0], c100 = g[1], c010 = g[sy], c110 = g[sy + 1];
  float c001 = g[sz], c101 = g[sz + 1], c011 = g[sz + sy], c111 = g[sz + sy + 1];
  float c00 = fmaf(fx, c100 - c000, c000), c10 = fmaf(fx, c110 - c010, c010);
  float c01 = fmaf(fx, c101 - c001, c001), c11 = fmaf(fx, c111 - c011, c011);
  float c0 = fmaf(fy, c10 - c00, c00), c1 = fmaf(fy, c11 - c01, c01);
  d = fmaf(fz, c1 - c0, c0);
  // d/dx: interpolate the x-differences in y,z; likewise for y and z
  float dx00 = c100 - c000, dx10 = c110 - c010, dx01 = c101 - c001, dx11 = c111 - c011;
  float dx0 = fmaf(fy, dx10 - dx00, dx00), dx1 = fmaf(fy, dx11 - dx01, dx01);
  float dy0 = c10 - c00, dy1 = c11 - c01;
  nrm.x = fmaf(fz, dx1 - dx0, dx0) * C.g_inv_h;
  nrm.y = fmaf(fz, dy1 - dy0, dy0) * C.g_inv_h;
  nrm.z = (c1 - c0) * C.g_inv_h;
  return true;
}

// ------------------------------------------------------------------ RBF skin value (row A7), fp64
// f(x) = sum w_i |x-c_i|^3 + a0 + a.x ; returned value f/|grad f| (normalize) -- fp64 because the
// cubic terms cancel to ~1e-4 of their magnitude.
struct SkinEval { double s, f, ginv; D3 nh; };
// 1/sqrt(r2) and sqrt(r2) in fp64 from an fp32 rsqrt seed (rel. error 2^-22) + one fp64 Newton step:
// rel. error 1.5*(2^-22)^2 = 8.5e-14, far below the 1e-10 the RBF sums need; r2 == 0 gives r = 0 and a
// finite (huge) inverse that only ever multiplies an exact zero.
FLASH_HD double drsqrt_fast(double r2) {
#ifdef __CUDA_ARCH__
  double y = (double)rsqrt_fast(fmaxf((float)r2, 1e-30f));
  return y * fma(-0.5 * r2, y * y, 1.5);
#else
  return r2 > 0.0 ? 1.0 / sqrt(r2) : 1e15;
#endif
}
FLASH_HD double dsqrt_fast(double r2) {
#ifdef __CUDA_ARCH__
  return r2 * drsqrt_fast(r2);
#else
  return sqrt(r2);
#endif
}
FLASH_HD SkinEval skin_eval(const double* __restrict__ c, const double* __restrict__ w, int n, bool normalize, double px, double py, double pz) {
  FLASH_COUNT(skin64, 1);
  double f = w[n] + w[n + 1] * px + w[n + 2] * py + w[n + 3] * pz;
  double gx = w[n + 1], gy = w[n + 2], gz = w[n + 3];
FLASH_SKIN_LOOP
  for (int i = 0; i < n; ++i) {
    double ux = px - c[3 * i], uy = py - c[3 * i + 1], uz = pz - c[3 * i + 2];
    double r2 = fma(ux, ux, fma(uy, uy, uz * uz));
    double r = dsqrt_fast(r2);
    double wr = w[i] * r;
    f = fma(wr, r2, f);
    double k = 3.0 * wr;
    gx = fma(k, ux, gx); gy = fma(k, uy, gy); gz = fma(k, uz, gz);
  }
  SkinEval e;
  e.f = f;
  if (!normalize) { e.s = f; e.ginv = 1.0; e.nh = d3(0, 0, 0); return e; }
  double g2 = gx * gx + gy * gy + gz * gz;
#ifdef __CUDA_ARCH__
  double gi = drsqrt_fast(g2);
  gi = gi * fma(-0.5 * g2, gi * gi, 1.5);   // second Newton step: full fp64 accuracy for the value itself
#else
  double gi = 1.0 / sqrt(g2);
#endif
  e.ginv = gi; e.s = f * gi; e.nh = d3(gx * gi, gy * gi, gz * gi);
  return e;
}

// Certified fp32 interval [lo, hi] for the skin value of one point (pre-filter of the min over
// surfaces: the fp64 evaluation above is needed only where the skin can be the minimum).
// cw = fp32 side table {cx,cy,cz,3w}[n], {a0,ax,ay,az} (tabf_fill).  Error model: inputs are rounded
// to fp32 (centres: eps*|c|, with |c| <= |p| + r), every operation adds eps relative; each cubic term
// 3w r^3 therefore carries at most ~16 eps |3w| r^2 (r + |p|_1), each gradient term 3w r u at most
// ~12 eps |3w| r (r + |p|_1), and the n-term sums add n*eps of the absolute sum.  K below is that
// bound with a factor >2 of slack; eps = 2^-23.
struct SkinBound { float lo, hi; };
FLASH_HD SkinBound skin_bound_f32(const float* __restrict__ cw, int n, bool normalize, float px, float py, float pz) {
  const float p1 = fabsf(px) + fabsf(py) + fabsf(pz);
  float f3 = 0.f, gx = 0.f, gy = 0.f, gz = 0.f, A = 0.f, B = 0.f;   // f3 = 3 * sum w r^3
FLASH_SKIN_LOOP
  for (int i = 0; i < n; ++i) {
    const F4 cc = reinterpret_cast<const F4*>(cw)[i];   // 16-byte aligned: one LDS.128
    const float ux = px - cc.x, uy = py - cc.y, uz = pz - cc.z, w3 = cc.w;
    const float r2 = fmaf(ux, ux, fmaf(uy, uy, uz * uz));
#ifdef __CUDA_ARCH__
    const float r = r2 * rsqrt_fast(fmaxf(r2, 1e-30f));
#else
    const float r = sqrtf(r2);
#endif
    const float wr = w3 * r;
    f3 = fmaf(wr, r2, f3);
    gx = fmaf(wr, ux, gx); gy = fmaf(wr, uy, gy); gz = fmaf(wr, uz, gz);
    const float t = fabsf(wr) * (r + p1);
    B += t;
    A = fmaf(t, r, A);
  }
  const float* a = cw + 4 * n;
  const float third = 1.0f / 3.0f;
  const float f = fmaf(f3, third, fmaf(a[1], px, fmaf(a[2], py, fmaf(a[3], pz, a[0]))));
  const float K = 1.1920929e-7f * (float)(2 * n + 40);
  const float pa = fabsf(a[0]) + fabsf(a[1] * px) + fabsf(a[2] * py) + fabsf(a[3] * pz);
  const float Ef = K * (A * third + pa) + 1e-30f;
  SkinBound o;
  if (!normalize) { o.lo = f - Ef - fabsf(f) * 1e-6f; o.hi = f + Ef + fabsf(f) * 1e-6f; return o; }
  gx += a[1]; gy += a[2]; gz += a[3];
  const float gn = sqrtf(fmaf(gx, gx, fmaf(gy, gy, gz * gz)));
  const float Eg = 2.f * K * (B + fabsf(a[1]) + fabsf(a[2]) + fabsf(a[3])) + 1e-30f;
  const float g_lo = gn * (1.f - 1e-6f) - Eg, g_hi = gn * (1.f + 1e-6f) + Eg;
  const float f_lo = f - Ef, f_hi = f + Ef;
  o.lo = -3.0e38f; o.hi = 3.0e38f;
  if (g_lo > 0.f) {
    if (f_lo > 0.f) { o.lo = (f_lo / g_hi) * (1.f - 1e-6f); o.hi = (f_hi / g_lo) * (1.f + 1e-6f); }
    else if (f_hi < 0.f) { o.lo = (f_lo / g_lo) * (1.f + 1e-6f); o.hi = (f_hi / g_hi) * (1.f - 1e-6f); }
    else { o.lo = (f_lo / g_lo) * (1.f + 1e-6f) - 1e-30f; o.hi = (f_hi / g_lo) * (1.f + 1e-6f) + 1e-30f; }
  }
  return o;
}
// Certified lower bound of the skin value over the WHOLE ball |x - c| <= rho (group-level cull: a skin whose
// bound exceeds the group's hull upper bound is the minimum for no point of the group, so the per-point
// pre-filter above is skipped for all 32 points).  Third-order Taylor expansion at the ball centre with the
// actual Hessian: for phi = r^3, grad = 3 r u, Hess = 3 (r I + u u^T / r), |D^3 phi [d,d,d]| <= 6 |d|^3 and
// |D^3 phi [d,d,.]| <= 6 |d|^2, so with W3 = sum |3 w_i|
//   f(x)      >= f(c) - |g| rho - |H|_F rho^2 / 2 - W3 rho^3 / 3
//   |grad f(x)| <= |g| + |H|_F rho + W3 rho^2
// and s = f / |grad f| >= f_lo / g_hi wherever f_lo > 0.  fp32 evaluation errors follow skin_bound_f32's model
// (centre = an fp32 point); the n-term error sums are replaced by n * (largest term).  Returns -3e38 when nothing
// can be certified.  One thread per ball (the cost kernel runs it once per group of its tile before the point
// loop: inside the loop the extra code costs more instruction fetch than the skipped pre-filters save).
FLASH_HD float skin_ball_lower_f32(const float* __restrict__ cw, int n, bool normalize, float cx, float cy, float cz, float rho) {
  float v[11];   // f3, g[3], Hxx, Hyy, Hzz, Hxy, Hxz, Hyz, sum |k|
  float m1 = 0.f, m2 = 0.f, m3 = 0.f;   // max |k|(r+p1), .. * r, .. * r^2
#pragma unroll
  for (int q = 0; q < 11; ++q) v[q] = 0.f;
  const float p1 = fabsf(cx) + fabsf(cy) + fabsf(cz);
  for (int i = 0; i < n; ++i) {
    const F4 cc = reinterpret_cast<const F4*>(cw)[i];
    const float ux = cx - cc.x, uy = cy - cc.y, uz = cz - cc.z, k = cc.w;
    const float r2 = fmaf(ux, ux, fmaf(uy, uy, uz * uz));
#ifdef __CUDA_ARCH__
    const float ri = rsqrt_fast(fmaxf(r2, 1e-30f));
#else
    const float ri = 1.0f / sqrtf(fmaxf(r2, 1e-30f));
#endif
    const float r = r2 * ri, kr = k * r, ki = k * ri;
    v[0] = fmaf(kr, r2, v[0]);
    v[1] = fmaf(kr, ux, v[1]); v[2] = fmaf(kr, uy, v[2]); v[3] = fmaf(kr, uz, v[3]);
    const float kx = ki * ux, ky = ki * uy, kz = ki * uz;
    v[4] += fmaf(kx, ux, kr); v[5] += fmaf(ky, uy, kr); v[6] += fmaf(kz, uz, kr);
    v[7] = fmaf(kx, uy, v[7]); v[8] = fmaf(kx, uz, v[8]); v[9] = fmaf(ky, uz, v[9]);
    const float t = fabsf(k) * (r + p1);
    v[10] += fabsf(k); m1 = fmaxf(m1, t); m2 = fmaxf(m2, t * r); m3 = fmaxf(m3, t * r2);
  }
  const float* a4 = cw + 4 * n;
  const float third = 1.0f / 3.0f, nf = (float)n;
  const float f = fmaf(v[0], third, fmaf(a4[1], cx, fmaf(a4[2], cy, fmaf(a4[3], cz, a4[0]))));
  const float K = 1.1920929e-7f * (float)(2 * n + 40);
  const float pa = fabsf(a4[0]) + fabsf(a4[1] * cx) + fabsf(a4[2] * cy) + fabsf(a4[3] * cz);
  const float Ef = K * (nf * m3 * third + pa) + 1e-30f;
  const float gx = v[1] + a4[1], gy = v[2] + a4[2], gz = v[3] + a4[3];
  const float gn = sqrtf(fmaf(gx, gx, fmaf(gy, gy, gz * gz))) * (1.f + 1e-6f) + 2.f * K * (nf * m2 + fabsf(a4[1]) + fabsf(a4[2]) + fabsf(a4[3]));
  const float hd = fmaf(v[4], v[4], fmaf(v[5], v[5], v[6] * v[6])), ho = fmaf(v[7], v[7], fmaf(v[8], v[8], v[9] * v[9]));
  const float hn = sqrtf(fmaf(2.f, ho, hd)) * (1.f + 1e-6f) + 8.f * K * nf * m1;   // Frobenius norm + evaluation error
  const float W3 = v[10] * (1.f + 1e-5f);   // >= sum |3 w_i|
  const float rh = rho * (1.f + 1e-6f);
  const float f_lo = (f - Ef - fabsf(f) * 1e-6f) - (rh * (gn + rh * (0.5f * hn + rh * third * W3))) * (1.f + 1e-5f);
  if (!(f_lo > 0.f)) return -3.0e38f;
  if (!normalize) return f_lo;
  const float g_hi = (gn + rh * (hn + rh * W3)) * (1.f + 1e-5f) + 1e-30f;
  return (f_lo / g_hi) * (1.f - 1e-6f);
}
// fp32 side table of one config: hull transforms (12 floats each) and the skin tables above
FLASH_HD void tabf_fill(const HotModel& H, const double* __restrict__ tab, float* __restrict__ tabf, int tid, int nthreads) {
  for (int i = tid; i < 12 * H.n_convex; i += nthreads) tabf[i] = (float)tab[i];
  for (int k = 0; k < H.n_skins; ++k) {
    const SkinC& K = H.skin[k];
    const double* c = tab + K.tab_off;
    const double* w = c + 3 * K.npts;
    float* dst = tabf + K.tabf_off;
    for (int e = tid; e < 4 * (K.npts + 1); e += nthreads) {
      const int i = e >> 2, a = e & 3;
      dst[e] = i < K.npts ? (a < 3 ? (float)c[3 * i + a] : (float)(3.0 * w[i])) : (float)w[K.npts + a];
    }
  }
}

// Adjoint contributions of ONE point (where this skin is the minimum) to centre i:
// out[0] = d(s^2)/d w_i, out[1..3] = explicit d(s^2)/d c_i.  i == n is the polynomial pseudo-centre:
// out[0] = d/d a0, out[1..3] = d/d a.   (derivation: DESIGN.md "RBF adjoint")
FLASH_HD void skin_adjoint_term(const double* __restrict__ c, const double* __restrict__ w, int n, bool normalize, int i,
                                double px, double py, double pz, double s, double ginv, D3 nh, double out[4]) {
  const double beta = normalize ? 2.0 * s * ginv : 2.0 * s;  // raw mode: s holds f
  const double sn = normalize ? s : 0.0;
  if (i == n) {
    out[0] = beta;
    out[1] = beta * (px - sn * nh.x); out[2] = beta * (py - sn * nh.y); out[3] = beta * (pz - sn * nh.z);
    return;
  }
  double ux = px - c[3 * i], uy = py - c[3 * i + 1], uz = pz - c[3 * i + 2];
  double r2 = ux * ux + uy * uy + uz * uz;
  double ri = drsqrt_fast(r2);
  double r = r2 * ri;
  double un = ux * nh.x + uy * nh.y + uz * nh.z;
  out[0] = beta * (r * r2 - sn * 3.0 * r * un);
  double wi = w[i];
  double k1 = -3.0 * wi * r;                              // d f / d c_i          = k1 * u
  double k2 = sn * wi * 3.0 * r;                          // -s * n.d(grad f)/dc_i = k2 * nh + k3 * u
  double k3 = r2 > 0.0 ? sn * wi * 3.0 * un * ri : 0.0;
  out[1] = beta * ((k1 + k3) * ux + k2 * nh.x);
  out[2] = beta * ((k1 + k3) * uy + k2 * nh.y);
  out[3] = beta * ((k1 + k3) * uz + k2 * nh.z);
}

// ------------------------------------------------------------------ min over surfaces (row A6)
struct GeoView {            // geometry blob (shared memory in the main kernel)
  const F4* f4;             // plane pairs, vertex pairs
  const uint32_t* u32;      // per-plane start triangle
  const float* grid_pool;   // grid mode
};
FLASH_HD HullView hull_view(const GeoView& G, const ConvexC& C) {
  HullView V;
  V.planes = G.f4 + C.off_planes; V.ptri = G.u32 + C.off_tris; V.va = G.f4 + C.off_verts;
  V.vz = reinterpret_cast<const float*>(G.f4 + C.off_vz); V.np = C.nplanes_pad; V.nv = C.nverts_pad;
  V.voff = reinterpret_cast<const unsigned short*>(G.u32 + C.off_voff);
  V.vlist = reinterpret_cast<const unsigned char*>(G.u32 + C.off_vlist);
  V.pgrp = reinterpret_cast<const unsigned short*>(G.u32 + C.off_pgrp);
  V.bn = G.f4 + C.off_bn; V.bk = G.u32 + C.off_bk; V.bp = G.f4 + C.off_bp; V.bv = G.f4 + C.off_bv; V.be = G.f4 + C.off_be; V.bsp_on = C.bsp_on;
  return V;
}

struct PointEval {
  float d;        // min over surfaces
  int arg;        // surface index of the first minimum (src/Flash.jl:267)
  int slot;       // cvx[] or skin[] slot of the winner
  int is_skin;    // winner is an interpolating skin
  F3 y, n;        // winner convex: point and distance gradient in the geometry frame
  SkinEval sk;    // winner skin
};

FLASH_HD bool lex_less(float d, int s, float best, int arg) { return d < best || (d == best && s < arg); }

// Bounds of hull j's signed distance over the ball B(centre, rho) of world points (the 32 Morton-
// neighbouring points one warp evaluates together).  The distance is 1-Lipschitz, so
//   d(y) >= aabb_lower_bound(centre) - rho   and   d(y) <= |centre - hull centroid| - inradius + rho.
// T is the fp32 copy of the world->geometry transform (error < 2e-6 m for |x| < 4 m; both bounds carry
// that margin), the geometry origin is the hull centroid.
FLASH_HD void hull_ball_bounds(const ConvexC& C, const float* __restrict__ T, float cx, float cy, float cz, float rho, float& lb, float& ub) {
  const F3 y = to_geometry_f32(T, cx, cy, cz);
  const float r = rho * (1.f + 1e-6f);
  lb = aabb_lower_bound(C.lo, C.hi, C.depth_bound, y) - r - 4e-6f;
  ub = sqrtf(dot(y, y)) * (1.f + 4e-6f) + 8e-6f - C.r_in + r;
}
// Hulls that can be the minimum for SOME point of the ball: lb_j <= min_k ub_k.  A culled hull is
// strictly farther than hull argmin(ub) at every point of the ball, so neither the min nor the argmin
// (first minimum) changes.  `first` = the most promising candidate (smallest lower bound).
FLASH_HD unsigned group_candidates(const HotModel& M, const float* __restrict__ tabf, float cx, float cy, float cz, float rho, int& first, float* min_ub_out = nullptr) {
  float lb[MAX_CONVEX];
  float min_ub = 3.0e38f;
  for (int j = 0; j < M.n_convex; ++j) {
    float ub;
    hull_ball_bounds(M.cvx[j], tabf + 12 * j, cx, cy, cz, rho, lb[j], ub);
    min_ub = fminf(min_ub, ub);
  }
  if (min_ub_out) *min_ub_out = min_ub;
  unsigned mask = 0; float lb0 = 3.0e38f; first = -1;
  for (int j = 0; j < M.n_convex; ++j)
    if (lb[j] <= min_ub) { mask |= 1u << j; if (lb[j] < lb0) { lb0 = lb[j]; first = j; } }
  return mask;
}

// Skins that are the minimum for NO point of the ball (bit k set = skip skin k): their certified lower bound over
// the ball exceeds min_ub, the smallest hull upper bound (group_candidates), by more than the fp32 rounding of
// the final comparison.  `cut` receives an upper bound of the group's minimum to start the hull scans from.
FLASH_HD float skin_cull_threshold(float min_ub) { return min_ub * (min_ub > 0.f ? 1.f + 2e-6f : 1.f - 2e-6f) + 2e-6f; }
FLASH_HD unsigned group_skin_cull(const HotModel& M, const float* __restrict__ tabf, float cx, float cy, float cz, float rho, float min_ub, float& cut) {
  unsigned skip = 0u;
  const float thr = skin_cull_threshold(min_ub);
  for (int k = 0; k < M.n_skins; ++k) {
    const SkinC& K = M.skin[k];
    if (skin_ball_lower_f32(tabf + K.tabf_off, K.npts, M.normalize != 0, cx, cy, cz, rho) > thr) skip |= 1u << k;
  }
  cut = skip ? thr : 3.0e38f;
  return skip;
}

// The whole group pre-pass of cost_kernel for one group (ball c, rho) of one configuration:
//   cand / first   hulls that can be the minimum for some point of the ball, the most promising one
//   feat           certified closest feature of `first` for the whole ball (hull_ball_feature), 0 if none
//   skip / cut     skins that are the minimum for no point of the ball; upper bound of the group's minimum
// The exact distance of `first` at the ball centre (a by-product of the feature search) tightens the upper bound
// of the group's minimum from |c - centroid| - inradius + rho to d(c) + rho, which culls more hulls and skins.
struct GroupPre { unsigned cand; int first; unsigned feat; unsigned skip; float cut; };
FLASH_HD GroupPre group_prepass(const HotModel& M, const GeoView& G, const float* __restrict__ tabf, float cx, float cy, float cz, float rho,
                                bool want_feat, bool want_skin_cull) {
  GroupPre o; o.cand = 0u; o.first = -1; o.feat = 0u; o.skip = 0u; o.cut = 3.0e38f;
  const int nc = M.n_convex;
  float min_ub = 3.0e38f, lb0 = 3.0e38f;
  float lb[MAX_CONVEX];
  for (int j = 0; j < nc; ++j) {
    float ubv;
    hull_ball_bounds(M.cvx[j], tabf + 12 * j, cx, cy, cz, rho, lb[j], ubv);
    min_ub = fminf(min_ub, ubv);
    if (lb[j] < lb0) { lb0 = lb[j]; o.first = j; }
  }
  if (want_feat && o.first >= 0) {
    const ConvexC& C = M.cvx[o.first];
    float dc;
    o.feat = hull_ball_feature(hull_view(G, C), to_geometry_f32(tabf + 12 * o.first, cx, cy, cz), rho, dc);
    const float ub2 = dc + rho * (1.f + 1e-5f) + 1e-5f;   // 1-Lipschitz: d(x) <= d(c) + rho
    if (dc < 3.0e37f) min_ub = fminf(min_ub, ub2);
  }
  for (int j = 0; j < nc; ++j)
    if (lb[j] <= min_ub || j == o.first) o.cand |= 1u << j;
  if (o.cand == 0u) { o.cand = nc >= 32 ? 0xffffffffu : ((1u << nc) - 1u); o.first = -1; o.feat = 0u; return o; }   // only with non-finite bounds
  if (want_skin_cull && M.n_skins > 0 && M.n_skins <= 8) o.skip = group_skin_cull(M, tabf, cx, cy, cz, rho, min_ub, o.cut);
  return o;
}

// Evaluates skin(x) = min over all surfaces for one world point under one configuration.
// `cand` = bit mask of the convex slots to consider (group_candidates, or all ones), `first` = the
// slot to evaluate first (-1: chosen here from the per-point bounds).  Within the candidates a body is
// evaluated exactly only if its per-point AABB lower bound can still beat the best value so far,
// which cannot change the min or the argmin.
// The point is passed in fp64: the cost path feeds fp32 points (exact in fp64), the ray marcher its
// fp64 sample positions.
FLASH_HD void eval_point(const HotModel& M, const GeoView& G, const double* __restrict__ tab, const float* __restrict__ tabf,
                         double px, double py, double pz, unsigned cand, int first, bool use_grid, bool fp32_point, PointEval& out,
                         unsigned skin_skip = 0u, float cut0 = 3.0e38f, unsigned feat = 0u) {
  const ConvexC* __restrict__ cvx = M.cvx;
  float best = 3.0e38f; int arg = 0x7fffffff; int slot = -1; int is_skin = 0;
  out.y = f3(0, 0, 0); out.n = f3(0, 0, 0);
  out.sk.s = 0; out.sk.f = 0; out.sk.ginv = 0; out.sk.nh = d3(0, 0, 0);
  const int nc = M.n_convex;
  const float fx = (float)px, fy = (float)py, fz = (float)pz;
  // Skins.  With convex bodies in the model the fp64 evaluation is deferred: a certified fp32 interval
  // [lo, hi] of each skin's value is enough to (a) give the hull scans a cutoff (hi) and (b) decide
  // afterwards whether the skin can still be the minimum (lo <= best hull); only then is it evaluated
  // in fp64.  The point must be fp32-representable for the interval to be valid (cost path); the ray
  // marcher passes fp64 sample positions and evaluates the skins directly.
  const bool defer = nc > 0 && fp32_point;
  float s_lo[MAX_SKINS];
  float cut = cut0;   // an upper bound of the final minimum that no surface index is attached to
  for (int k = 0; k < M.n_skins; ++k) {
    const SkinC& K = M.skin[k];
    if (defer && ((skin_skip >> k) & 1u)) { s_lo[k] = 3.0e38f; continue; }   // group_skin_cull: never the minimum here
    if (defer) {
      const SkinBound sb = skin_bound_f32(tabf + K.tabf_off, K.npts, M.normalize != 0, fx, fy, fz);
      s_lo[k] = sb.lo;
      cut = fminf(cut, sb.hi);
    } else {
      const double* c = tab + K.tab_off;
      SkinEval e = skin_eval(c, c + 3 * K.npts, K.npts, M.normalize != 0, px, py, pz);
      float d = (float)e.s;
      if (lex_less(d, K.surface, best, arg)) { best = d; arg = K.surface; slot = k; is_skin = 1; out.sk = e; }
    }
  }
  FLASH_COUNT(points, 1);
  if (first < 0) {   // no group information: the most promising body by the per-point bound
    float lb0 = 3.0e38f;
    for (int j = 0; j < nc; ++j) {
      if (!((cand >> j) & 1u)) continue;
      FLASH_COUNT(lb_tests, 1);
      const float l = aabb_lower_bound(cvx[j].lo, cvx[j].hi, cvx[j].depth_bound, to_geometry_f32(tabf + 12 * j, fx, fy, fz));
      if (l < lb0) { lb0 = l; first = j; }
    }
  }
  // `first`, then the remaining candidates in surface order
  const bool single = (cand & (cand - 1u)) == 0u;
  unsigned rest = first >= 0 ? (cand & ~(1u << first)) : cand;
  int j = first;
  while (true) {
    if (j < 0) {
      if (!rest) break;
#ifdef __CUDA_ARCH__
      j = __ffs(rest) - 1;
#else
      j = __builtin_ctz(rest);
#endif
      rest &= rest - 1u;
    }
    const ConvexC& C = cvx[j];
    // Per-point use of the group's certified feature / GJK warm start.  Compiled into the device loop only on request
    // (-DFLASH_FEAT_PATH): measured on K5 it LOSES -- 54.9 ms with it, 49.0 with the code present but never taken,
    // 42.2 ms without the code (round 2, B200; K2 0.110 / 0.088 / 0.080 ms, K4 4.65 / 4.41 / 3.97 ms): the point loop is
    // instruction-fetch and register bound (80 registers: 460 B of spills with the path, 216 B without).  The host
    // replay (tests/hostcheck) keeps it, so the certificates stay tested.
#if !defined(__CUDA_ARCH__) || defined(FLASH_FEAT_PATH)
    if ((feat & 3u) != 0u && j == first) {
      const F3 y = to_geometry(tab + 12 * j, px, py, pz);
      F3 n; float d; bool ok = true;
      if (feat & FEAT_CERT) {
        // the group's ball lies in the Voronoi region of one feature of this hull (hull_ball_feature): closed form
        FLASH_COUNT(feat_evals, 1);
        d = hull_feature_eval(hull_view(G, C), feat, y, n);
      } else ok = hull_sd_warm(hull_view(G, C), feat, y, d, n);   // GJK from the simplex found at the ball centre
      if (ok) {
        if (lex_less(d, C.surface, best, arg)) { best = d; arg = C.surface; slot = j; is_skin = 0; out.y = y; out.n = n; }
        j = -1;
        continue;
      }
    }
#endif
    bool skip = false;
    if (!single) {
      FLASH_COUNT(lb_tests, 1);
      const float l = aabb_lower_bound(C.lo, C.hi, C.depth_bound, to_geometry_f32(tabf + 12 * j, fx, fy, fz));
      skip = l > best || (l == best && C.surface > arg) || l > cut;
    }
    if (!skip) {
      F3 y = to_geometry(tab + 12 * j, px, py, pz);
      F3 n; float d;
      bool done = false;
      if (use_grid && C.g_on) done = grid_sd(C, G.grid_pool, y, d, n);
#if defined(FLASH_COUNTERS) && !defined(__CUDA_ARCH__)
      const long long narrow0 = flashdev::counters().narrow;
#endif
#ifdef FLASH_GJK_PATH
      if (!done) d = hull_sd(hull_view(G, C), y, fminf(best, cut), n);
#else
      if (!done) { FLASH_COUNT(narrow, 1); d = hull_sd_bsp(hull_view(G, C), y, n); }
#endif
#if defined(FLASH_COUNTERS) && !defined(__CUDA_ARCH__)
      if (flashdev::counters().narrow != narrow0) flashdev::counters().exact_mask |= 1u << j;
#endif
      if (lex_less(d, C.surface, best, arg)) { best = d; arg = C.surface; slot = j; is_skin = 0; out.y = y; out.n = n; }
    }
    j = -1;
  }
  if (defer) {
    for (int k = 0; k < M.n_skins; ++k) {
      if (s_lo[k] > best) continue;   // certified: this skin is strictly above the best hull
      const SkinC& K = M.skin[k];
      const double* c = tab + K.tab_off;
      SkinEval e = skin_eval(c, c + 3 * K.npts, K.npts, M.normalize != 0, px, py, pz);
      float d = (float)e.s;
      if (lex_less(d, K.surface, best, arg)) { best = d; arg = K.surface; slot = k; is_skin = 1; out.sk = e; }
    }
  }
  out.d = best; out.arg = arg; out.slot = slot; out.is_skin = is_skin;
}

// ------------------------------------------------------------------ epilogue (rows A2-A5 gradient)
// acc: this config's accumulators after the sum over all points (and ranks).  Produces the raw cost
// and d cost / d x, including the in-cost quaternion normalisation (src/gradientdescent.jl:30) and
// the deformation regulariser (:33-37).  Serial; a few thousand flops per config.
FLASH_HD void epilogue_config(const ModelC& M, const double* x, const double* tab, const double* aux, const double* acc,
                              double w_def, double* scratch /* 6*n_bodies + MAX_SKIN_M */, double* cost, double* grad) {
  const int nb = M.n_bodies;
  double* F = scratch;            // [nb][3]
  double* Mo = scratch + 3 * nb;  // [nb][3]
  double* lam = scratch + 6 * nb; // [M]
  for (int i = 0; i < 6 * nb; ++i) scratch[i] = 0;
  if (grad) for (int k = 0; k < M.nx; ++k) grad[k] = 0;
  double c = acc[0];
  if (grad) {
    // convex wrenches, accumulated in the geometry frame: f = sum g, tau = sum y x g, g = -2 d n
    for (int j = 0; j < M.n_convex; ++j) {
      const ConvexC& C = M.cvx[j];
      if (C.body < 0) continue;
      double S[12];
      xf_compose(aux + aux_X(M, C.body), C.frame, S);
      const double* a = acc + 1 + 6 * j;
      D3 f = xf_rot(S, d3(a[0], a[1], a[2]));
      D3 m = xf_rot(S, d3(a[3], a[4], a[5])) + cross(xf_t(S), f);
      F[3 * C.body] += f.x; F[3 * C.body + 1] += f.y; F[3 * C.body + 2] += f.z;
      Mo[3 * C.body] += m.x; Mo[3 * C.body + 1] += m.y; Mo[3 * C.body + 2] += m.z;
    }
    for (int k = 0; k < M.n_skins; ++k) {
      const SkinC& K = M.skin[k];
      const int n = K.npts, Mm = K.M;
      const double* cc = tab + K.tab_off;
      const double* w = cc + 3 * n;
      const double* gY = acc + K.acc_off;
      const double* gC = gY + n + 4;
      const double* LU = aux + aux_skin0(M) + K.aux_off;
      const double* piv = LU + Mm * Mm;
      for (int i = 0; i < Mm; ++i) lam[i] = gY[i];
      lu_solve(LU, piv, lam, Mm);   // symmetric system: A^-T gY = A^-1 gY
      D3 la = Mm == n + 4 ? d3(lam[n + 1], lam[n + 2], lam[n + 3]) : d3(0, 0, 0);
      D3 av = d3(w[n + 1], w[n + 2], w[n + 3]);
      for (int i = 0; i < n; ++i) {
        D3 ci = d3(cc[3 * i], cc[3 * i + 1], cc[3 * i + 2]);
        D3 Gi = d3(gC[3 * i], gC[3 * i + 1], gC[3 * i + 2]);
        for (int j = 0; j < n; ++j) {
          D3 u = ci - d3(cc[3 * j], cc[3 * j + 1], cc[3 * j + 2]);
          double r = sqrt(dot(u, u));
          Gi = Gi - u * (3.0 * r * (lam[i] * w[j] + lam[j] * w[i]));
        }
        Gi = Gi - av * lam[i] - la * w[i];
        int b = M.skin_pt_body[K.pt_off + i];
        if (b >= 0) {
          D3 m = cross(ci, Gi);
          F[3 * b] += Gi.x; F[3 * b + 1] += Gi.y; F[3 * b + 2] += Gi.z;
          Mo[3 * b] += m.x; Mo[3 * b + 1] += m.y; Mo[3 * b + 2] += m.z;
        }
        if (K.kind == SURF_DEFORMABLE_SKIN && i < K.nsurf) {
          D3 gd = b >= 0 ? xf_rotT(aux + aux_X(M, b), Gi) : Gi;
          double* o = grad + M.nq + K.deform_off + 3 * i;
          o[0] += gd.x; o[1] += gd.y; o[2] += gd.z;
        }
      }
    }
    for (int b = nb - 1; b >= 0; --b) {  // subtree wrenches: children follow parents
      int p = M.body[b].parent;
      if (p >= 0) for (int k = 0; k < 3; ++k) { F[3 * p + k] += F[3 * b + k]; Mo[3 * p + k] += Mo[3 * b + k]; }
    }
    for (int b = 0; b < nb; ++b) {
      const BodyC& B = M.body[b];
      if (B.jtype == JOINT_FIXED) continue;
      const double* P = aux + aux_P(M, b);
      D3 Fb = d3(F[3 * b], F[3 * b + 1], F[3 * b + 2]), Mb = d3(Mo[3 * b], Mo[3 * b + 1], Mo[3 * b + 2]);
      if (B.jtype == JOINT_REVOLUTE) {
        D3 wv = xf_rot(P, d3(B.axis[0], B.axis[1], B.axis[2]));
        grad[B.qoff] += dot(wv, Mb - cross(xf_t(P), Fb));
      } else {
        const double* Xb = aux + aux_X(M, b);
        D3 Gp = xf_rotT(P, Mb - cross(xf_t(Xb), Fb));
        D3 Fp = xf_rotT(P, Fb);
        const double nrm = aux[aux_qnorm(M, b)];
        const double* q = x + B.qoff;
        const double qw = q[0] / nrm;
        D3 v = d3(q[1] / nrm, q[2] / nrm, q[3] / nrm);
        D3 gv = Gp * qw + cross(Gp, v);
        const double inv = 2.0 / nrm;
        grad[B.qoff] += -dot(Gp, v) * inv;
        grad[B.qoff + 1] += gv.x * inv; grad[B.qoff + 2] += gv.y * inv; grad[B.qoff + 3] += gv.z * inv;
        grad[B.qoff + 4] += Fp.x; grad[B.qoff + 5] += Fp.y; grad[B.qoff + 6] += Fp.z;
      }
    }
  }
  double reg = 0;
  for (int k = M.nq; k < M.nx; ++k) { reg += x[k] * x[k]; if (grad) grad[k] += 2.0 * w_def * x[k]; }
  *cost = c + w_def * reg;
}

}  // namespace flashdev
