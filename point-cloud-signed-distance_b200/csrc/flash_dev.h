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
struct Counters { long long points, narrow, lb_tests, skin64, bsp_levels; unsigned exact_mask; };   // exact_mask: hulls whose narrow phase ran for the LAST point
inline Counters& counters() { static Counters c = {0, 0, 0, 0, 0, 0u}; return c; }
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
  // closest-feature BSP tree (flash_bsp.h) in the geometry blob: node planes / plane table / vertices / edges (F4
  // indices), child words (u32 index); node count
  int off_bn, off_bk, off_bp, off_bv, off_be, n_nodes;
  float lo[3], hi[3];                        // AABB of the vertices, geometry frame
  float depth_bound, r_in;                   // >= max penetration depth; <= inradius about the geometry origin
  double frame[12];                          // geometry frame -> body frame
  // grid mode (body-frame SDF grid): origin, spacing, dims, offset into the grid pool (floats)
  float g_org[3]; float g_inv_h; int g_dim[3]; int g_on; long long g_off;
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
// Signed distance from y (geometry frame) to a convex hull through its closest-feature BSP tree (flash_bsp.h):
// one plane test per level locates the cell of y, whose feature -- a face, an edge or a vertex -- gives the distance
// and its gradient in closed form.  Inside the hull the cell's face is the plane of largest (least negative)
// distance: the exact depth.  (EnhancedGJK `gjk!` is the reference algorithm, src/Flash.jl:238-243; outside it
// converges to the same number, inside its value is iteration-path dependent -- DESIGN.md 2.)
// Every lane of a warp runs the same ~11 instructions per level whatever its path; round 1's per-point GJK (plane
// hint, start triangle, support queries, simplex updates) was 45 % of cost_kernel's instructions at 17-23 of 32
// active lanes.
//
// Per hull in the geometry blob: node planes F4 {n, -o} + child words u32 (lo | hi << 16; s = n.y - o > 0 -> hi);
// a child with bit 15 set is a leaf: 0x8000 | kind << 13 | index, kind 0 = face (index into the plane table
// F4 {n, -o}), 1 = vertex (vertex table F4 {x, y, z, 0}), 2 = edge (edge table F4 {unit direction, bits(first vertex)}).
struct HullView { const F4* bn; const uint32_t* bk; const F4* bp; const F4* bv; const F4* be; };

// descend from `start` (a node index, or already a leaf word) to the leaf of y
FLASH_HD unsigned bsp_locate(const HullView& V, unsigned start, F3 y) {
  unsigned k = start;
  while (!(k & 0x8000u)) {
    FLASH_COUNT(bsp_levels, 1);
#ifdef __CUDA_ARCH__
    const float4 pp = __ldg(reinterpret_cast<const float4*>(V.bn) + k);
    F4 p; p.x = pp.x; p.y = pp.y; p.z = pp.z; p.w = pp.w;
    const uint32_t w = __ldg(V.bk + k);
#else
    const F4 p = V.bn[k];
    const uint32_t w = V.bk[k];
#endif
    const float s = fmaf(p.x, y.x, fmaf(p.y, y.y, fmaf(p.z, y.z, p.w)));
    k = s > 0.f ? (w >> 16) : (w & 0xffffu);
  }
  return k;
}
// Descend with a whole ball B(yc, r): as long as the ball lies on one side of the node's plane every point of it
// takes the same child.  Returns the first node the ball straddles (or the leaf that contains it): the start word
// from which the points of the ball finish their own descent (cost_kernel's group records).  A point within
// rounding distance of a plane may end up on either side: the distance is continuous across every cell boundary.
FLASH_HD unsigned bsp_descend_ball(const HullView& V, F3 yc, float r) {
  unsigned k = 0u;
  while (!(k & 0x8000u)) {
    const F4 p = V.bn[k];
    const uint32_t w = V.bk[k];
    const float s = fmaf(p.x, yc.x, fmaf(p.y, yc.y, fmaf(p.z, yc.z, p.w)));
    if (s > r) k = w >> 16; else if (s < -r) k = w & 0xffffu; else break;
  }
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
FLASH_HD float hull_sd(const HullView& V, F3 y, F3& nrm, unsigned start = 0u) {
  FLASH_COUNT(narrow, 1);
  return bsp_feature_sd(V, bsp_locate(V, start, y), y, nrm);
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
struct GeoView {            // geometry blob (global memory, read through L1)
  const F4* f4;             // BSP node planes, plane / vertex / edge tables
  const uint32_t* u32;      // BSP child words
  const float* grid_pool;   // grid mode
};
FLASH_HD HullView hull_view(const GeoView& G, const ConvexC& C) {
  HullView V;
  V.bn = G.f4 + C.off_bn; V.bk = G.u32 + C.off_bk; V.bp = G.f4 + C.off_bp; V.bv = G.f4 + C.off_bv; V.be = G.f4 + C.off_be;
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

// The group pre-pass of cost_kernel (group_kernel: one thread per group of 32 points and configuration; ball c, rho):
//   cand / first   hulls that can be the minimum for some point of the ball, the nearest one at the centre
//   skip / cut     skins that are the minimum for no point of the ball; upper bound of the group's minimum
//   starts         per candidate, in evaluation order (first, then the others by index; three at most): the BSP node
//                  from which the points of the ball finish their descent (bsp_descend_ball), or already their leaf
// Bounds: the cheap AABB / inradius bounds of hull_ball_bounds first; with `exact`, every hull they keep is evaluated
// exactly at the ball centre (one BSP descent), and the distance being 1-Lipschitz, d_j(c) -+ rho bounds it over
// the ball far more tightly (K5: 3.0 candidates per group from the AABB bounds).  Hulls with lb_j > min_k ub_k are
// strictly farther than hull argmin(ub) at every point of the ball, so neither the min nor the argmin changes.
// Not with grid mode, whose per-point values deviate from the exact distance by the grid error.
struct GroupPre { unsigned cand; int first; unsigned skip; float cut; unsigned long long starts; };
FLASH_HD GroupPre group_prepass(const HotModel& M, const GeoView& G, const float* __restrict__ tabf, float cx, float cy, float cz, float rho,
                                bool exact, bool want_skin_cull) {
  GroupPre o; o.cand = 0u; o.first = -1; o.skip = 0u; o.cut = 3.0e38f; o.starts = 0ull;
  const int nc = M.n_convex;
  float min_ub = 3.0e38f;
  float lb[MAX_CONVEX];
  unsigned short st[MAX_CONVEX];
  for (int j = 0; j < nc; ++j) {
    float ubv;
    hull_ball_bounds(M.cvx[j], tabf + 12 * j, cx, cy, cz, rho, lb[j], ubv);
    min_ub = fminf(min_ub, ubv);
    st[j] = 0;
  }
  if (exact) {
    const float r = rho * (1.f + 1e-5f) + 1e-5f;   // ball radius + slack for the fp32 transform of the centre
    // the hull with the smallest AABB bound first: its exact distance usually rules out all the others
    int j0 = 0;
    for (int j = 1; j < nc; ++j) if (lb[j] < lb[j0]) j0 = j;
    for (int t = 0; t < nc; ++t) {
      const int j = t == 0 ? j0 : (t <= j0 ? t - 1 : t);
      if (lb[j] > min_ub) continue;
      const HullView V = hull_view(G, M.cvx[j]);
      const F3 yc = to_geometry_f32(tabf + 12 * j, cx, cy, cz);
      const unsigned s0 = bsp_descend_ball(V, yc, r);
      F3 nn;
      const float dc = bsp_feature_sd(V, bsp_locate(V, s0, yc), yc, nn);
      st[j] = (unsigned short)s0;
      lb[j] = fmaxf(lb[j], dc - r);
      min_ub = fminf(min_ub, dc + r);
    }
  }
  float lb0 = 3.0e38f;
  for (int j = 0; j < nc; ++j)
    if (lb[j] <= min_ub) { o.cand |= 1u << j; if (lb[j] < lb0) { lb0 = lb[j]; o.first = j; } }
  if (o.cand == 0u) { o.cand = nc >= 32 ? 0xffffffffu : ((1u << nc) - 1u); o.first = -1; return o; }   // only with non-finite bounds
  {
    int t = 1;
    o.starts = (unsigned long long)st[o.first];
    for (int j = 0; j < nc && t < 4; ++j)
      if (j != o.first && ((o.cand >> j) & 1u)) { o.starts |= (unsigned long long)st[j] << (16 * t); ++t; }
  }
  if (want_skin_cull && M.n_skins > 0 && M.n_skins <= 8) o.skip = group_skin_cull(M, tabf, cx, cy, cz, rho, min_ub, o.cut);
  return o;
}

// Evaluates skin(x) = min over all surfaces for one world point under one configuration.
// `cand` = bit mask of the convex slots to consider (group_prepass, or all ones), `first` = the
// slot to evaluate first (-1: chosen here from the per-point bounds), `starts` = the BSP start words of the
// candidates in evaluation order, 16 bits each (0 = the root).  Within the candidates a body is evaluated
// exactly only if its per-point AABB lower bound can still beat the best value so far, which cannot change
// the min or the argmin.
// The point is passed in fp64: the cost path feeds fp32 points (exact in fp64), the ray marcher its
// fp64 sample positions.
FLASH_HD void eval_point(const HotModel& M, const GeoView& G, const double* __restrict__ tab, const float* __restrict__ tabf,
                         double px, double py, double pz, unsigned cand, int first, bool use_grid, bool fp32_point, PointEval& out,
                         unsigned skin_skip = 0u, float cut0 = 3.0e38f, unsigned long long starts = 0ull) {
  const ConvexC* __restrict__ cvx = M.cvx;
  float best = 3.0e38f; int arg = 0x7fffffff; int slot = -1; int is_skin = 0;
  out.y = f3(0, 0, 0); out.n = f3(0, 0, 0);
  out.sk.s = 0; out.sk.f = 0; out.sk.ginv = 0; out.sk.nh = d3(0, 0, 0);
  const int nc = M.n_convex;
  const float fx = (float)px, fy = (float)py, fz = (float)pz;
  // Skins.  With convex bodies in the model the fp64 evaluation is deferred: a certified fp32 interval
  // [lo, hi] of each skin's value is enough to (a) give the hull tests a cutoff (hi) and (b) decide
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
    starts = 0ull;
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
    const unsigned start = (unsigned)(starts & 0xffffull);
    starts >>= 16;
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
      if (!done) flashdev::counters().exact_mask |= 1u << j;
#endif
      if (!done) d = hull_sd(hull_view(G, C), y, n, start);
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
