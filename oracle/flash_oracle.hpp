// flash_oracle.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU fp64 restatement of the tracking hot path of the Julia package `Flash`
// (JuliaTagBot/point-cloud-signed-distance).  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may use anything under oracle/.
//
// PARITY STATUS (see DESIGN.md "Oracle"):
//   * RBF skin: pinned by the reference's single known-answer test, test/runtests.jl:17
//     (beanbag skin(100,0,0) ~ 99.0 rtol 2e-2; this restatement gives 98.8930).
//   * convex signed distance (EnhancedGJK @404de6a), kinematics (RigidBodyDynamics 0.0.2),
//     gradient (ForwardDiff), optimiser step (SimpleGradientDescent @0fcc1f9): PARITY UNPINNED.
//     Those packages are pinned in REQUIRE.dev:18-28 but their source is not in the reference
//     tree and Julia is not installed; the published algorithms are restated and anchored on the
//     reference's call sites (cited at each function).
//
// Everything is templated on the scalar T (double or Dual<N>) because the reference obtains its
// gradient by pushing ForwardDiff duals through the whole pipeline
// (examples/irb_and_squishable.ipynb:474-489, src/gradientdescent.jl:49-57).
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/flashsdf.h"

namespace fo {

// ---------------------------------------------------------------- forward-mode dual numbers
template <int N>
struct Dual {
  double v;
  double d[N];
  Dual() : v(0) { for (int i = 0; i < N; ++i) d[i] = 0; }
  Dual(double a) : v(a) { for (int i = 0; i < N; ++i) d[i] = 0; }
};
template <int N> inline Dual<N> operator+(const Dual<N>& a, const Dual<N>& b) { Dual<N> r; r.v = a.v + b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] + b.d[i]; return r; }
template <int N> inline Dual<N> operator-(const Dual<N>& a, const Dual<N>& b) { Dual<N> r; r.v = a.v - b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] - b.d[i]; return r; }
template <int N> inline Dual<N> operator-(const Dual<N>& a) { Dual<N> r; r.v = -a.v; for (int i = 0; i < N; ++i) r.d[i] = -a.d[i]; return r; }
template <int N> inline Dual<N> operator*(const Dual<N>& a, const Dual<N>& b) { Dual<N> r; r.v = a.v * b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
template <int N> inline Dual<N> operator/(const Dual<N>& a, const Dual<N>& b) { Dual<N> r; double ib = 1.0 / b.v; r.v = a.v * ib; for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * ib; return r; }
template <int N> inline Dual<N> operator+(const Dual<N>& a, double b) { Dual<N> r = a; r.v += b; return r; }
template <int N> inline Dual<N> operator+(double b, const Dual<N>& a) { Dual<N> r = a; r.v += b; return r; }
template <int N> inline Dual<N> operator-(const Dual<N>& a, double b) { Dual<N> r = a; r.v -= b; return r; }
template <int N> inline Dual<N> operator-(double b, const Dual<N>& a) { Dual<N> r = -a; r.v += b; return r; }
template <int N> inline Dual<N> operator*(const Dual<N>& a, double b) { Dual<N> r; r.v = a.v * b; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b; return r; }
template <int N> inline Dual<N> operator*(double b, const Dual<N>& a) { return a * b; }
template <int N> inline Dual<N> operator/(const Dual<N>& a, double b) { return a * (1.0 / b); }
template <int N> inline Dual<N> operator/(double a, const Dual<N>& b) { return Dual<N>(a) / b; }
template <int N> inline Dual<N>& operator+=(Dual<N>& a, const Dual<N>& b) { a = a + b; return a; }
template <int N> inline Dual<N>& operator-=(Dual<N>& a, const Dual<N>& b) { a = a - b; return a; }
template <int N> inline Dual<N> sqrt(const Dual<N>& a) { Dual<N> r; r.v = std::sqrt(a.v); double h = 0.5 / r.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * h; return r; }
template <int N> inline Dual<N> sin(const Dual<N>& a) { Dual<N> r; r.v = std::sin(a.v); double c = std::cos(a.v); for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * c; return r; }
template <int N> inline Dual<N> cos(const Dual<N>& a) { Dual<N> r; r.v = std::cos(a.v); double s = -std::sin(a.v); for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * s; return r; }
inline double sqrt(double a) { return std::sqrt(a); }
inline double sin(double a) { return std::sin(a); }
inline double cos(double a) { return std::cos(a); }
inline double val(double a) { return a; }
template <int N> inline double val(const Dual<N>& a) { return a.v; }

// ---------------------------------------------------------------- small vector algebra
template <class T> struct V3 { T x, y, z; };
typedef V3<double> V3d;
template <class T> inline V3<T> operator+(const V3<T>& a, const V3<T>& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class T> inline V3<T> operator-(const V3<T>& a, const V3<T>& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class T, class S> inline V3<T> scale(const V3<T>& a, const S& s) { return {a.x * s, a.y * s, a.z * s}; }
template <class T> inline T dot(const V3<T>& a, const V3<T>& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <class T> inline V3<T> cross(const V3<T>& a, const V3<T>& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
template <class T> inline T norm(const V3<T>& a) { return sqrt(dot(a, a)); }
template <class T> inline V3d val(const V3<T>& a) { return {val(a.x), val(a.y), val(a.z)}; }
template <class T> inline V3<T> lift(const V3d& a) { return {T(a.x), T(a.y), T(a.z)}; }

// p_out = R p + t
template <class T> struct Xf {
  T R[9];
  V3<T> t;
  static Xf identity() { Xf X; for (int i = 0; i < 9; ++i) X.R[i] = T(i % 4 == 0 ? 1.0 : 0.0); X.t = {T(0.0), T(0.0), T(0.0)}; return X; }
  V3<T> rot(const V3<T>& p) const { return {R[0] * p.x + R[1] * p.y + R[2] * p.z, R[3] * p.x + R[4] * p.y + R[5] * p.z, R[6] * p.x + R[7] * p.y + R[8] * p.z}; }
  V3<T> apply(const V3<T>& p) const { return rot(p) + t; }
  V3<T> inv_apply(const V3<T>& p) const { V3<T> u = p - t; return {R[0] * u.x + R[3] * u.y + R[6] * u.z, R[1] * u.x + R[4] * u.y + R[7] * u.z, R[2] * u.x + R[5] * u.y + R[8] * u.z}; }
};
template <class T> inline Xf<T> compose(const Xf<T>& A, const Xf<T>& B) {  // A*B
  Xf<T> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.R[3 * i + j] = A.R[3 * i] * B.R[j] + A.R[3 * i + 1] * B.R[3 + j] + A.R[3 * i + 2] * B.R[6 + j];
  C.t = A.apply(B.t);
  return C;
}
template <class T> inline Xf<T> xf_from12(const double* m) {
  Xf<T> X;
  for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) X.R[3 * i + j] = T(m[4 * i + j]); }
  X.t = {T(m[3]), T(m[7]), T(m[11])};
  return X;
}

// ---------------------------------------------------------------- model
struct Body { int parent, jtype, qoff; double axis[3]; double j2p[12]; };

struct Hull {                      // convex hull of a vertex set, surface frame
  std::vector<V3d> verts;          // welded input vertices
  std::vector<std::array<double, 4>> planes;  // unique supporting planes: n.x - o <= 0 inside
  std::vector<std::array<int, 3>> tris;       // fan triangulation of every face polygon
};

struct Surface {
  int kind;
  int body;              // convex
  double frame[12];      // convex
  Hull hull;             // convex
  std::vector<V3d> pts;  // skins: control points in their body frame, surface first then skeleton
  std::vector<int> pt_body;
  int n_surface;         // skins: number of value-0 points
  int deform_offset;     // deformable: index of first deformation scalar within deformation_data
};

struct Model {
  std::vector<Body> bodies;
  std::vector<Surface> surfaces;
  int nq = 0, ndef = 0;  // ndef = number of deformable points (3 scalars each)
  int normalize_by_gradient = 1;
  int poly_degree = 1;
  int nx() const { return nq + 3 * ndef; }  // src/Flash.jl:90
};

// Brute-force convex hull (O(n^4), n <= a few hundred): every vertex triple whose plane has all
// points on one side is a supporting plane; coplanar triples are merged into one polygon.
inline void build_hull(const std::vector<V3d>& in, Hull& h) {
  // weld duplicates (STL stores 3 vertices per triangle, SURVEY 7 "STL vertex duplication")
  double diam = 0;
  for (auto& a : in) for (auto& b : in) diam = std::max(diam, norm(a - b));
  const double weld = 1e-9 * std::max(diam, 1e-30);
  for (auto& p : in) {
    bool dup = false;
    for (auto& q : h.verts) if (norm(p - q) <= weld) { dup = true; break; }
    if (!dup) h.verts.push_back(p);
  }
  const int n = (int)h.verts.size();
  if (n < 4) throw std::runtime_error("convex surface needs >= 4 non-coplanar vertices");
  const double eps = 1e-9 * diam;
  std::vector<std::array<double, 4>> planes;
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j)
      for (int k = j + 1; k < n; ++k) {
        V3d nn = cross(h.verts[j] - h.verts[i], h.verts[k] - h.verts[i]);
        double l = norm(nn);
        if (l < 1e-12 * diam * diam) continue;
        nn = scale(nn, 1.0 / l);
        double o = dot(nn, h.verts[i]);
        double mx = -1e300, mn = 1e300;
        for (int m = 0; m < n; ++m) { double s = dot(nn, h.verts[m]) - o; mx = std::max(mx, s); mn = std::min(mn, s); }
        if (mx > eps && mn < -eps) continue;
        if (mx > eps) { nn = scale(nn, -1.0); o = -o; }   // flip so that all points are on the <= 0 side
        bool seen = false;
        for (auto& p : planes) {
          if (std::fabs(p[0] - nn.x) + std::fabs(p[1] - nn.y) + std::fabs(p[2] - nn.z) < 1e-7 && std::fabs(p[3] - o) < 1e-7 * std::max(diam, 1.0)) { seen = true; break; }
        }
        if (!seen) planes.push_back({nn.x, nn.y, nn.z, o});
      }
  if (planes.size() < 4) throw std::runtime_error("degenerate convex surface");
  h.planes = planes;
  // face polygons -> triangles
  for (auto& p : planes) {
    V3d nn{p[0], p[1], p[2]};
    std::vector<int> on;
    for (int m = 0; m < n; ++m) if (std::fabs(dot(nn, h.verts[m]) - p[3]) <= 10 * eps + 1e-7 * diam * 1e-2) on.push_back(m);
    if (on.size() < 3) continue;
    // corners of the face polygon: 2-D convex hull (monotone chain) of the vertices on the plane, so that vertices
    // inside the face or on its edges (legal input, not corners) cannot spoil the fan
    V3d e1 = std::fabs(nn.x) < 0.9 ? cross(nn, V3d{1, 0, 0}) : cross(nn, V3d{0, 1, 0});
    e1 = scale(e1, 1.0 / norm(e1));
    V3d e2 = cross(nn, e1);
    std::vector<std::array<double, 3>> q;
    for (int m : on) q.push_back({dot(h.verts[m], e1), dot(h.verts[m], e2), (double)m});
    std::sort(q.begin(), q.end());
    auto turn = [](const std::array<double, 3>& a, const std::array<double, 3>& b, const std::array<double, 3>& c) { return (b[0] - a[0]) * (c[1] - a[1]) - (b[1] - a[1]) * (c[0] - a[0]); };
    std::vector<std::array<double, 3>> ch;
    for (int pass = 0; pass < 2; ++pass) {
      const size_t start = ch.size();
      for (size_t i = 0; i < q.size(); ++i) {
        const auto& pt = pass == 0 ? q[i] : q[q.size() - 1 - i];
        while (ch.size() >= start + 2 && turn(ch[ch.size() - 2], ch[ch.size() - 1], pt) <= 1e-12 * diam * diam) ch.pop_back();
        ch.push_back(pt);
      }
      ch.pop_back();
    }
    for (size_t t = 1; t + 1 < ch.size(); ++t) h.tris.push_back({(int)ch[0][2], (int)ch[t][2], (int)ch[t + 1][2]});
  }
}

inline Model model_from_desc(const flash_model_desc* d, int normalize_by_gradient, int poly_degree) {
  Model m;
  m.normalize_by_gradient = normalize_by_gradient;
  m.poly_degree = poly_degree;
  int qoff = 0;
  for (int b = 0; b < d->n_bodies; ++b) {
    Body B;
    B.parent = d->body_parent[b];
    if (B.parent >= b) throw std::runtime_error("bodies must be in tree order");
    B.jtype = d->joint_type[b];
    std::memcpy(B.axis, d->joint_axis + 3 * b, sizeof B.axis);
    std::memcpy(B.j2p, d->joint_to_parent + 12 * b, sizeof B.j2p);
    B.qoff = qoff;
    qoff += B.jtype == FLASH_JOINT_REVOLUTE ? 1 : B.jtype == FLASH_JOINT_QUAT_FLOATING ? 7 : 0;
    m.bodies.push_back(B);
  }
  m.nq = qoff;
  int doff = 0;
  for (int s = 0; s < d->n_surfaces; ++s) {
    Surface S;
    S.kind = d->surface_kind[s];
    S.body = d->surface_body[s];
    std::memcpy(S.frame, d->surface_frame + 12 * s, sizeof S.frame);
    S.n_surface = d->n_surface_points[s];
    S.deform_offset = -1;
    std::vector<V3d> pts;
    for (int p = d->point_offset[s]; p < d->point_offset[s + 1]; ++p) {
      pts.push_back({d->points[3 * p], d->points[3 * p + 1], d->points[3 * p + 2]});
      S.pt_body.push_back(d->point_body ? d->point_body[p] : -1);
    }
    if (S.kind == FLASH_SURF_CONVEX) build_hull(pts, S.hull);
    else S.pts = pts;
    if (S.kind == FLASH_SURF_DEFORMABLE_SKIN) { S.deform_offset = doff; doff += 3 * S.n_surface; }
    m.surfaces.push_back(S);
  }
  m.ndef = doff / 3;
  return m;
}

// ---------------------------------------------------------------- kinematics
// Standard tree forward kinematics (RigidBodyDynamics 0.0.2 `transform_to_root`; call sites
// src/Flash.jl:147,154,165,193,248).  q is already normalised (src/gradientdescent.jl:30).
template <class T> inline Xf<T> joint_transform(const Body& B, const T* q) {
  Xf<T> J = Xf<T>::identity();
  if (B.jtype == FLASH_JOINT_REVOLUTE) {  // Revolute(axis): rotation by q about axis (src/models.jl:26)
    double ax = B.axis[0], ay = B.axis[1], az = B.axis[2];
    double l = std::sqrt(ax * ax + ay * ay + az * az); ax /= l; ay /= l; az /= l;
    T c = cos(q[B.qoff]), s = sin(q[B.qoff]);
    T C = 1.0 - c;
    J.R[0] = c + C * (ax * ax);      J.R[1] = C * (ax * ay) - s * az; J.R[2] = C * (ax * az) + s * ay;
    J.R[3] = C * (ay * ax) + s * az; J.R[4] = c + C * (ay * ay);      J.R[5] = C * (ay * az) - s * ax;
    J.R[6] = C * (az * ax) - s * ay; J.R[7] = C * (az * ay) + s * ax; J.R[8] = c + C * (az * az);
  } else if (B.jtype == FLASH_JOINT_QUAT_FLOATING) {  // q = [w x y z tx ty tz] (src/models.jl:79,106)
    T w = q[B.qoff], x = q[B.qoff + 1], y = q[B.qoff + 2], z = q[B.qoff + 3];
    J.R[0] = 1.0 - 2.0 * (y * y + z * z); J.R[1] = 2.0 * (x * y - w * z);       J.R[2] = 2.0 * (x * z + w * y);
    J.R[3] = 2.0 * (x * y + w * z);       J.R[4] = 1.0 - 2.0 * (x * x + z * z); J.R[5] = 2.0 * (y * z - w * x);
    J.R[6] = 2.0 * (x * z - w * y);       J.R[7] = 2.0 * (y * z + w * x);       J.R[8] = 1.0 - 2.0 * (x * x + y * y);
    J.t = {q[B.qoff + 4], q[B.qoff + 5], q[B.qoff + 6]};
  }
  return J;
}

// normalize! for QuaternionFloating joints, src/gradientdescent.jl:19-26 (inside the differentiated cost)
template <class T> inline void normalize_quats(const Model& m, std::vector<T>& q) {
  for (auto& B : m.bodies)
    if (B.jtype == FLASH_JOINT_QUAT_FLOATING) {
      T n = sqrt(q[B.qoff] * q[B.qoff] + q[B.qoff + 1] * q[B.qoff + 1] + q[B.qoff + 2] * q[B.qoff + 2] + q[B.qoff + 3] * q[B.qoff + 3]);
      for (int k = 0; k < 4; ++k) q[B.qoff + k] = q[B.qoff + k] / n;
    }
}

template <class T> inline void forward_kinematics(const Model& m, const T* q, std::vector<Xf<T>>& X) {
  X.resize(m.bodies.size());
  for (size_t b = 0; b < m.bodies.size(); ++b) {
    const Body& B = m.bodies[b];
    Xf<T> local = compose(xf_from12<T>(B.j2p), joint_transform<T>(B, q));
    X[b] = B.parent < 0 ? local : compose(X[B.parent], local);
  }
}

// ---------------------------------------------------------------- RBF skin
// `InterpolatingSurface(points, values, XCubed(), true)` (src/Flash.jl:207-213), restated as the
// Turk & O'Brien variational implicit surface with phi(r)=r^3 and a linear polynomial; the value
// returned is f/|grad f| when normalize_by_gradient (the only r^3 definition that
// satisfies test/runtests.jl:17).
template <class T> struct SkinField {
  std::vector<V3<T>> c;
  std::vector<T> w;
  T a0;
  V3<T> a;
  bool normalize;
  T eval(const V3<T>& x) const {
    T f = a0 + dot(a, x);
    V3<T> g = a;
    for (size_t i = 0; i < c.size(); ++i) {
      V3<T> u = x - c[i];
      T r2 = dot(u, u);
      T r = val(r2) > 0 ? sqrt(r2) : T(0.0);
      f = f + w[i] * (r * r2);
      T k = 3.0 * w[i] * r;
      g = g + scale(u, k);
    }
    if (!normalize) return f;
    return f / norm(g);
  }
};

template <class T> inline void lu_solve(std::vector<T>& A, std::vector<T>& b, int n) {
  for (int k = 0; k < n; ++k) {
    int piv = k; double best = std::fabs(val(A[k * n + k]));
    for (int i = k + 1; i < n; ++i) if (std::fabs(val(A[i * n + k])) > best) { best = std::fabs(val(A[i * n + k])); piv = i; }
    if (best == 0) throw std::runtime_error("singular RBF system");
    if (piv != k) { for (int j = 0; j < n; ++j) std::swap(A[k * n + j], A[piv * n + j]); std::swap(b[k], b[piv]); }
    for (int i = k + 1; i < n; ++i) {
      T f = A[i * n + k] / A[k * n + k];
      for (int j = k + 1; j < n; ++j) A[i * n + j] = A[i * n + j] - f * A[k * n + j];
      b[i] = b[i] - f * b[k];
    }
  }
  for (int i = n - 1; i >= 0; --i) {
    T s = b[i];
    for (int j = i + 1; j < n; ++j) s = s - A[i * n + j] * b[j];
    b[i] = s / A[i * n + i];
  }
}

// surface_points / skeleton_points (src/Flash.jl:152-205) + skin (:207-213)
template <class T> inline SkinField<T> build_skin(const Model& m, const Surface& S, const std::vector<Xf<T>>& X, const T* deform) {
  SkinField<T> F;
  const int n = (int)S.pts.size();
  F.normalize = m.normalize_by_gradient != 0;
  for (int i = 0; i < n; ++i) {
    V3<T> p = lift<T>(S.pts[i]);
    if (S.kind == FLASH_SURF_DEFORMABLE_SKIN && i < S.n_surface) {  // point + deformation, in the body frame (:176-181)
      const T* dd = deform + S.deform_offset + 3 * i;
      p = p + V3<T>{dd[0], dd[1], dd[2]};
    }
    F.c.push_back(S.pt_body[i] < 0 ? p : X[S.pt_body[i]].apply(p));
  }
  const int np = m.poly_degree >= 1 ? 4 : 1;
  const int M = n + np;
  std::vector<T> A((size_t)M * M, T(0.0)), b(M, T(0.0));
  for (int i = 0; i < n; ++i) {
    for (int j = 0; j < n; ++j) {
      if (i == j) continue;
      V3<T> u = F.c[i] - F.c[j];
      T r2 = dot(u, u);
      A[i * M + j] = val(r2) > 0 ? r2 * sqrt(r2) : T(0.0);
    }
    A[i * M + n] = T(1.0); A[n * M + i] = T(1.0);
    if (np == 4) {
      A[i * M + n + 1] = F.c[i].x; A[i * M + n + 2] = F.c[i].y; A[i * M + n + 3] = F.c[i].z;
      A[(n + 1) * M + i] = F.c[i].x; A[(n + 2) * M + i] = F.c[i].y; A[(n + 3) * M + i] = F.c[i].z;
    }
    b[i] = T(i < S.n_surface ? 0.0 : -1.0);  // zeros(surface) , -1 + zeros(skeleton)  (:211)
  }
  lu_solve(A, b, M);
  F.w.assign(b.begin(), b.begin() + n);
  F.a0 = b[n];
  F.a = np == 4 ? V3<T>{b[n + 1], b[n + 2], b[n + 3]} : V3<T>{T(0.0), T(0.0), T(0.0)};
  return F;
}

// ---------------------------------------------------------------- convex signed distance
// Closest point on triangle abc to p (Ericson, Real-Time Collision Detection 5.1.5); comparisons on
// values so that the same code differentiates the active region's projection formula.
template <class T> inline V3<T> closest_on_triangle(const V3<T>& p, const V3<T>& a, const V3<T>& b, const V3<T>& c) {
  V3<T> ab = b - a, ac = c - a, ap = p - a;
  T d1 = dot(ab, ap), d2 = dot(ac, ap);
  if (val(d1) <= 0 && val(d2) <= 0) return a;
  V3<T> bp = p - b;
  T d3 = dot(ab, bp), d4 = dot(ac, bp);
  if (val(d3) >= 0 && val(d4) <= val(d3)) return b;
  T vc = d1 * d4 - d3 * d2;
  if (val(vc) <= 0 && val(d1) >= 0 && val(d3) <= 0) { T v = d1 / (d1 - d3); return a + scale(ab, v); }
  V3<T> cp = p - c;
  T d5 = dot(ab, cp), d6 = dot(ac, cp);
  if (val(d6) >= 0 && val(d5) <= val(d6)) return c;
  T vb = d5 * d2 - d1 * d6;
  if (val(vb) <= 0 && val(d2) >= 0 && val(d6) <= 0) { T w = d2 / (d2 - d6); return a + scale(ac, w); }
  T va = d3 * d6 - d5 * d4;
  if (val(va) <= 0 && val(d4 - d3) >= 0 && val(d5 - d6) >= 0) { T w = (d4 - d3) / ((d4 - d3) + (d5 - d6)); return b + scale(c - b, w); }
  T denom = 1.0 / (va + vb + vc);
  T v = vb * denom, w = vc * denom;
  return a + scale(ab, v) + scale(ac, w);
}
template <class T> inline V3<T> closest_on_segment(const V3<T>& p, const V3<T>& a, const V3<T>& b) {
  V3<T> ab = b - a;
  T t = dot(p - a, ab) / dot(ab, ab);
  if (val(t) <= 0) return a;
  if (val(t) >= 1) return b;
  return a + scale(ab, t);
}

// Ericson's closest point on a triangle, double only, also returning barycentric weights whose
// exact zeros identify the supporting sub-simplex.
inline V3d tri_closest_bary(const V3d& p, const V3d& a, const V3d& b, const V3d& c, double w[3]) {
  V3d ab = b - a, ac = c - a, ap = p - a;
  double d1 = dot(ab, ap), d2 = dot(ac, ap);
  if (d1 <= 0 && d2 <= 0) { w[0] = 1; w[1] = 0; w[2] = 0; return a; }
  V3d bp = p - b;
  double d3 = dot(ab, bp), d4 = dot(ac, bp);
  if (d3 >= 0 && d4 <= d3) { w[0] = 0; w[1] = 1; w[2] = 0; return b; }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) { double v = d1 / (d1 - d3); w[0] = 1 - v; w[1] = v; w[2] = 0; return a + scale(ab, v); }
  V3d cp = p - c;
  double d5 = dot(ab, cp), d6 = dot(ac, cp);
  if (d6 >= 0 && d5 <= d6) { w[0] = 0; w[1] = 0; w[2] = 1; return c; }
  double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) { double t = d2 / (d2 - d6); w[0] = 1 - t; w[1] = 0; w[2] = t; return a + scale(ac, t); }
  double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) { double t = (d4 - d3) / ((d4 - d3) + (d5 - d6)); w[0] = 0; w[1] = 1 - t; w[2] = t; return b + scale(c - b, t); }
  double denom = 1.0 / (va + vb + vc);
  double v = vb * denom, t = vc * denom;
  w[0] = 1 - v - t; w[1] = v; w[2] = t;
  return a + scale(ab, v) + scale(ac, t);
}

// GJK, point vs convex vertex set, fp64 (the reference's algorithm for this row: EnhancedGJK
// `gjk!`, src/Flash.jl:238-243; restated from Gilbert-Johnson-Keerthi with Ericson's simplex
// sub-distance; max_iter 100 as SURVEY 8a row A8).  Returns the <=3 vertices supporting the
// closest point (0 if y turned out to be inside).
inline int gjk_closest_feature(const Hull& h, const V3d& y, int start, int out[3]) {
  const int nv = (int)h.verts.size();
  int sx[4]; int ns = 1; sx[0] = start;
  V3d q = h.verts[start];
  for (int iter = 0; iter < 100; ++iter) {
    V3d dir = y - q;
    double dd = dot(dir, dir);
    if (dd == 0) break;
    int best = -1; double bs = -1e300;
    for (int v = 0; v < nv; ++v) { double s = dot(dir, h.verts[v]); if (s > bs) { bs = s; best = v; } }
    bool in = false;
    for (int i = 0; i < ns; ++i) if (sx[i] == best) in = true;
    if (in || bs - dot(dir, q) <= 1e-15 * std::max(1.0, dd)) break;
    sx[ns++] = best;
    if (ns == 2) {
      V3d a = h.verts[sx[0]], b = h.verts[sx[1]];
      double t = dot(y - a, b - a) / dot(b - a, b - a);
      if (t <= 0) { ns = 1; q = a; } else if (t >= 1) { sx[0] = sx[1]; ns = 1; q = b; } else q = a + scale(b - a, t);
      continue;
    }
    static const int F4[4][3] = {{0, 1, 2}, {0, 1, 3}, {0, 2, 3}, {1, 2, 3}};
    int nf = ns == 3 ? 1 : 4;
    if (ns == 4) {  // y inside the tetrahedron -> inside the hull
      V3d A = h.verts[sx[0]], B = h.verts[sx[1]], C = h.verts[sx[2]], D = h.verts[sx[3]];
      auto side = [&](const V3d& a, const V3d& b, const V3d& c, const V3d& d) { V3d nn = cross(b - a, c - a); return dot(nn, y - a) * dot(nn, d - a); };
      if (side(A, B, C, D) >= 0 && side(A, B, D, C) >= 0 && side(A, C, D, B) >= 0 && side(B, C, D, A) >= 0) return 0;
    }
    double bd = 1e300; int bf = 0; V3d bq = q; double bw[3] = {1, 0, 0};
    for (int f = 0; f < nf; ++f) {
      double w[3];
      V3d cq = tri_closest_bary(y, h.verts[sx[F4[f][0]]], h.verts[sx[F4[f][1]]], h.verts[sx[F4[f][2]]], w);
      double d2 = dot(y - cq, y - cq);
      if (d2 < bd) { bd = d2; bf = f; bq = cq; bw[0] = w[0]; bw[1] = w[1]; bw[2] = w[2]; }
    }
    int t3[3] = {sx[F4[bf][0]], sx[F4[bf][1]], sx[F4[bf][2]]};
    q = bq; ns = 0;
    for (int i = 0; i < 3; ++i) if (bw[i] > 0) sx[ns++] = t3[i];
  }
  for (int i = 0; i < ns && i < 3; ++i) out[i] = sx[i];
  return std::min(ns, 3);
}

enum { CONVEX_BRUTE = 0, CONVEX_GJK = 1 };

// Signed distance from world point x to the hull posed by `pose` (surface frame -> world).
//   outside: Euclidean distance to the hull (what gjk!(...).signed_distance returns, to GJK's atol);
//   inside : -(penetration depth) = max over supporting planes of the signed plane distance.
//            (EnhancedGJK returns the depth w.r.t. its terminal simplex there -- iteration-path
//            dependent and not reproducible without its source; SURVEY 7 hard parts.  Stated
//            deviation, identical in the CUDA path.)
template <class T> inline T convex_sd(const Hull& h, const Xf<T>& pose, const V3d& x, int mode) {
  V3<T> y = pose.inv_apply(lift<T>(x));
  V3d yv = val(y);
  int fmax = 0; double D = -1e300;
  for (size_t f = 0; f < h.planes.size(); ++f) {
    double s = h.planes[f][0] * yv.x + h.planes[f][1] * yv.y + h.planes[f][2] * yv.z - h.planes[f][3];
    if (s > D) { D = s; fmax = (int)f; }
  }
  if (D <= 0) return y.x * h.planes[fmax][0] + y.y * h.planes[fmax][1] + y.z * h.planes[fmax][2] - h.planes[fmax][3];
  if (mode == CONVEX_GJK) {
    int sup[3];
    int start = 0; double bs = -1e300;   // start from the vertex nearest to y
    for (size_t v = 0; v < h.verts.size(); ++v) { double s = -dot(yv - h.verts[v], yv - h.verts[v]); if (s > bs) { bs = s; start = (int)v; } }
    int ns = gjk_closest_feature(h, yv, start, sup);
    V3<T> q;
    if (ns == 0) return y.x * h.planes[fmax][0] + y.y * h.planes[fmax][1] + y.z * h.planes[fmax][2] - h.planes[fmax][3];
    if (ns == 1) q = lift<T>(h.verts[sup[0]]);
    else if (ns == 2) q = closest_on_segment(y, lift<T>(h.verts[sup[0]]), lift<T>(h.verts[sup[1]]));
    else q = closest_on_triangle(y, lift<T>(h.verts[sup[0]]), lift<T>(h.verts[sup[1]]), lift<T>(h.verts[sup[2]]));
    return norm(y - q);
  }
  int bt = 0; double bd = 1e300;
  for (size_t t = 0; t < h.tris.size(); ++t) {
    V3d cq = closest_on_triangle(yv, h.verts[h.tris[t][0]], h.verts[h.tris[t][1]], h.verts[h.tris[t][2]]);
    double d2 = dot(yv - cq, yv - cq);
    if (d2 < bd) { bd = d2; bt = (int)t; }
  }
  V3<T> q = closest_on_triangle(y, lift<T>(h.verts[h.tris[bt][0]]), lift<T>(h.verts[h.tris[bt][1]]), lift<T>(h.verts[h.tris[bt][2]]));
  return norm(y - q);
}

// ---------------------------------------------------------------- skin(state): min over surfaces
template <class T> struct PosedSurfaces {
  const Model* m;
  std::vector<SkinField<T>> skins;  // indexed by surface (empty for convex)
  std::vector<Xf<T>> poses;         // indexed by surface (convex)
  int convex_mode;
};

// unflatten! (src/gradientdescent.jl:13-17) + normalize! (:30) + surfaces(state) (src/Flash.jl:261-263)
template <class T> inline PosedSurfaces<T> pose_surfaces(const Model& m, const T* x, int convex_mode) {
  PosedSurfaces<T> P; P.m = &m; P.convex_mode = convex_mode;
  std::vector<T> q(x, x + m.nq);
  normalize_quats(m, q);
  std::vector<Xf<T>> X;
  forward_kinematics(m, q.data(), X);
  const T* deform = x + m.nq;
  P.skins.resize(m.surfaces.size()); P.poses.resize(m.surfaces.size());
  for (size_t s = 0; s < m.surfaces.size(); ++s) {
    const Surface& S = m.surfaces[s];
    if (S.kind == FLASH_SURF_CONVEX) {
      Xf<T> fr = xf_from12<T>(S.frame);
      P.poses[s] = S.body < 0 ? fr : compose(X[S.body], fr);   // transform_to_root(state, surface.frame) (:248)
    } else P.skins[s] = build_skin<T>(m, S, X, deform);
  }
  return P;
}

// x -> minimum(s(x) for s in all_surfaces) (src/Flash.jl:267); argmin = first minimum
template <class T> inline T skin_eval(const PosedSurfaces<T>& P, const V3d& x, int* argmin) {
  T best = T(0.0); int arg = -1;
  for (size_t s = 0; s < P.m->surfaces.size(); ++s) {
    const Surface& S = P.m->surfaces[s];
    T d = S.kind == FLASH_SURF_CONVEX ? convex_sd<T>(S.hull, P.poses[s], x, P.convex_mode) : P.skins[s].eval(lift<T>(x));
    if (arg < 0 || val(d) < val(best)) { best = d; arg = (int)s; }
  }
  if (argmin) *argmin = arg;
  return best;
}

// cost (src/gradientdescent.jl:28-39): sum_p skin(p)^2 + w * sum deformation.^2, RAW (not / N)
template <class T> inline T cost(const Model& m, const T* x, const double* pts, int64_t n, double w_def, int convex_mode, int32_t* argmin) {
  PosedSurfaces<T> P = pose_surfaces<T>(m, x, convex_mode);
  T c = T(0.0);
  // blocked accumulation (Base.mapreduce is pairwise; at fp64 the orders agree to ~1e-13)
  const int64_t B = 1024;
  for (int64_t i0 = 0; i0 < n; i0 += B) {
    T part = T(0.0);
    for (int64_t i = i0; i < std::min(n, i0 + B); ++i) {
      int a;
      T d = skin_eval<T>(P, V3d{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]}, &a);
      if (argmin) argmin[i] = a;
      part = part + d * d;
    }
    c = c + part;
  }
  T reg = T(0.0);
  for (int k = m.nq; k < m.nx(); ++k) reg = reg + x[k] * x[k];
  return c + w_def * reg;
}

// ForwardDiff's chunk size rule (pickchunksize, threshold 10): 63 -> 9, 43 -> 9
// (examples/irb_and_squishable.ipynb:479-483 shows Dual{9} for 63 states)
inline int pick_chunk(int n) {
  if (n <= 10) return n;
  int nchunks = (n + 9) / 10;
  return (n + nchunks - 1) / nchunks;
}

// ---------------------------------------------------------------- analytic (adjoint) gradient, fp64
// NOT how the reference computes its gradient (that is forward-mode duals, above); this is the
// single-pass algorithm the CUDA path uses, restated in fp64 so that (a) the dual gradient has an
// independent cross-check and (b) bench.py can time the best CPU algorithm beside the faithful one.
struct AdjointAcc {
  double cost = 0;
  std::vector<V3d> F, M;                  // per body: sum of forces G and moments z x G (world origin)
  std::vector<std::vector<double>> gY;    // per surface (skins): dC/dy explicit
  std::vector<std::vector<V3d>> gC;       // per surface (skins): dC/dc explicit
  void init(const Model& m) {
    F.assign(m.bodies.size(), V3d{0, 0, 0}); M = F;
    gY.resize(m.surfaces.size()); gC.resize(m.surfaces.size());
    for (size_t s = 0; s < m.surfaces.size(); ++s)
      if (m.surfaces[s].kind != FLASH_SURF_CONVEX) { gY[s].assign(m.surfaces[s].pts.size() + 4, 0.0); gC[s].assign(m.surfaces[s].pts.size(), V3d{0, 0, 0}); }
  }
  void add(const AdjointAcc& o) {
    cost += o.cost;
    for (size_t b = 0; b < F.size(); ++b) { F[b] = F[b] + o.F[b]; M[b] = M[b] + o.M[b]; }
    for (size_t s = 0; s < gY.size(); ++s) { for (size_t i = 0; i < gY[s].size(); ++i) gY[s][i] += o.gY[s][i]; for (size_t i = 0; i < gC[s].size(); ++i) gC[s][i] = gC[s][i] + o.gC[s][i]; }
  }
};

struct AdjointCtx {
  const Model* m;
  std::vector<double> qn;           // normalised q
  std::vector<double> qnorm;        // per body: |quat| before normalisation (1 for others)
  std::vector<Xf<double>> X;        // body -> world
  PosedSurfaces<double> P;
};

inline AdjointCtx adjoint_prepare(const Model& m, const double* x, int convex_mode) {
  AdjointCtx A; A.m = &m;
  A.qn.assign(x, x + m.nq); A.qnorm.assign(m.bodies.size(), 1.0);
  for (size_t b = 0; b < m.bodies.size(); ++b)
    if (m.bodies[b].jtype == FLASH_JOINT_QUAT_FLOATING) { const double* q = x + m.bodies[b].qoff; A.qnorm[b] = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]); }
  normalize_quats(m, A.qn);
  forward_kinematics(m, A.qn.data(), A.X);
  A.P = pose_surfaces<double>(m, x, convex_mode);
  return A;
}

// one point: value, argmin, and its contribution to the accumulators
inline void adjoint_point(const AdjointCtx& A, const V3d& x, AdjointAcc& acc, int* argmin) {
  const Model& m = *A.m;
  int arg; double d = skin_eval<double>(A.P, x, &arg);
  if (argmin) *argmin = arg;
  acc.cost += d * d;
  const Surface& S = m.surfaces[arg];
  if (S.kind == FLASH_SURF_CONVEX) {
    if (S.body < 0) return;
    // world gradient of the distance by re-deriving the closest feature
    const Xf<double>& pose = A.P.poses[arg];
    V3d y = pose.inv_apply(x);
    int fmax = 0; double D = -1e300;
    for (size_t f = 0; f < S.hull.planes.size(); ++f) { double v = S.hull.planes[f][0] * y.x + S.hull.planes[f][1] * y.y + S.hull.planes[f][2] * y.z - S.hull.planes[f][3]; if (v > D) { D = v; fmax = (int)f; } }
    V3d nb;
    if (D <= 0) nb = V3d{S.hull.planes[fmax][0], S.hull.planes[fmax][1], S.hull.planes[fmax][2]};
    else {
      int bt = 0; double bd = 1e300;
      for (size_t t = 0; t < S.hull.tris.size(); ++t) { V3d cq = closest_on_triangle(y, S.hull.verts[S.hull.tris[t][0]], S.hull.verts[S.hull.tris[t][1]], S.hull.verts[S.hull.tris[t][2]]); double d2 = dot(y - cq, y - cq); if (d2 < bd) { bd = d2; bt = (int)t; } }
      V3d q = closest_on_triangle(y, S.hull.verts[S.hull.tris[bt][0]], S.hull.verts[S.hull.tris[bt][1]], S.hull.verts[S.hull.tris[bt][2]]);
      nb = scale(y - q, 1.0 / norm(y - q));
    }
    V3d nw = pose.rot(nb);
    V3d G = scale(nw, -2.0 * d);
    acc.F[S.body] = acc.F[S.body] + G;
    acc.M[S.body] = acc.M[S.body] + cross(x, G);
    return;
  }
  const SkinField<double>& F = A.P.skins[arg];
  const int n = (int)F.c.size();
  double f = F.a0 + dot(F.a, x); V3d g = F.a;
  for (int i = 0; i < n; ++i) { V3d u = x - F.c[i]; double r = norm(u); f += F.w[i] * r * r * r; g = g + scale(u, 3.0 * F.w[i] * r); }
  double gn = norm(g);
  double beta, s; V3d nh{0, 0, 0};
  if (F.normalize) { s = f / gn; beta = 2.0 * s / gn; nh = scale(g, 1.0 / gn); } else { s = 0; beta = 2.0 * f; }
  std::vector<double>& gY = acc.gY[arg]; std::vector<V3d>& gC = acc.gC[arg];
  for (int i = 0; i < n; ++i) {
    V3d u = x - F.c[i]; double r = norm(u); double un = dot(u, nh);
    gY[i] += beta * (r * r * r - s * 3.0 * r * un);
    V3d t = scale(u, -3.0 * F.w[i] * r);
    if (F.normalize) { t = t + scale(nh, s * F.w[i] * 3.0 * r); if (r > 0) t = t + scale(u, s * F.w[i] * 3.0 * un / r); }
    gC[i] = gC[i] + scale(t, beta);
  }
  gY[n] += beta;
  V3d ga = scale(x - scale(nh, s), beta);
  gY[n + 1] += ga.x; gY[n + 2] += ga.y; gY[n + 3] += ga.z;
}

// epilogue: adjoint RBF solve, point forces -> joint-space gradient
inline void adjoint_finish(const AdjointCtx& A, AdjointAcc& acc, const double* x, double w_def, double* grad) {
  const Model& m = *A.m;
  const int nx = m.nx();
  for (int k = 0; k < nx; ++k) grad[k] = 0;
  for (size_t si = 0; si < m.surfaces.size(); ++si) {
    const Surface& S = m.surfaces[si];
    if (S.kind == FLASH_SURF_CONVEX) continue;
    const SkinField<double>& F = A.P.skins[si];
    const int n = (int)F.c.size();
    const int np = m.poly_degree >= 1 ? 4 : 1, M = n + np;
    std::vector<double> Am((size_t)M * M, 0.0), lam(M, 0.0);
    for (int i = 0; i < n; ++i) {
      for (int j = 0; j < n; ++j) { double r = norm(F.c[i] - F.c[j]); Am[i * M + j] = r * r * r; }
      Am[i * M + n] = 1; Am[n * M + i] = 1;
      if (np == 4) { Am[i * M + n + 1] = Am[(n + 1) * M + i] = F.c[i].x; Am[i * M + n + 2] = Am[(n + 2) * M + i] = F.c[i].y; Am[i * M + n + 3] = Am[(n + 3) * M + i] = F.c[i].z; }
    }
    for (int i = 0; i < M; ++i) lam[i] = acc.gY[si][i < n ? i : n + (i - n)];
    lu_solve(Am, lam, M);   // symmetric system: A^-T = A^-1
    V3d la = np == 4 ? V3d{lam[n + 1], lam[n + 2], lam[n + 3]} : V3d{0, 0, 0};
    for (int i = 0; i < n; ++i) {
      V3d G = acc.gC[si][i];
      for (int j = 0; j < n; ++j) { V3d u = F.c[i] - F.c[j]; double r = norm(u); G = G - scale(u, 3.0 * r * (lam[i] * F.w[j] + lam[j] * F.w[i])); }
      G = G - scale(F.a, lam[i]) - scale(la, F.w[i]);
      int b = S.pt_body[i];
      if (b >= 0) { acc.F[b] = acc.F[b] + G; acc.M[b] = acc.M[b] + cross(F.c[i], G); }
      if (S.kind == FLASH_SURF_DEFORMABLE_SKIN && i < S.n_surface) {
        V3d gd = b >= 0 ? V3d{A.X[b].R[0] * G.x + A.X[b].R[3] * G.y + A.X[b].R[6] * G.z, A.X[b].R[1] * G.x + A.X[b].R[4] * G.y + A.X[b].R[7] * G.z, A.X[b].R[2] * G.x + A.X[b].R[5] * G.y + A.X[b].R[8] * G.z} : G;
        double* o = grad + m.nq + S.deform_offset + 3 * i;
        o[0] += gd.x; o[1] += gd.y; o[2] += gd.z;
      }
    }
  }
  // subtree wrenches (children come after parents)
  for (int b = (int)m.bodies.size() - 1; b >= 0; --b) { int p = m.bodies[b].parent; if (p >= 0) { acc.F[p] = acc.F[p] + acc.F[b]; acc.M[p] = acc.M[p] + acc.M[b]; } }
  for (size_t b = 0; b < m.bodies.size(); ++b) {
    const Body& B = m.bodies[b];
    if (B.jtype == FLASH_JOINT_FIXED) continue;
    Xf<double> Pk = xf_from12<double>(B.j2p);
    if (B.parent >= 0) Pk = compose(A.X[B.parent], Pk);
    if (B.jtype == FLASH_JOINT_REVOLUTE) {
      double l = std::sqrt(B.axis[0] * B.axis[0] + B.axis[1] * B.axis[1] + B.axis[2] * B.axis[2]);
      V3d w = Pk.rot(V3d{B.axis[0] / l, B.axis[1] / l, B.axis[2] / l});
      grad[B.qoff] += dot(w, acc.M[b] - cross(Pk.t, acc.F[b]));
    } else {
      V3d Gw = acc.M[b] - cross(A.X[b].t, acc.F[b]);
      V3d Gp{Pk.R[0] * Gw.x + Pk.R[3] * Gw.y + Pk.R[6] * Gw.z, Pk.R[1] * Gw.x + Pk.R[4] * Gw.y + Pk.R[7] * Gw.z, Pk.R[2] * Gw.x + Pk.R[5] * Gw.y + Pk.R[8] * Gw.z};
      V3d Fp{Pk.R[0] * acc.F[b].x + Pk.R[3] * acc.F[b].y + Pk.R[6] * acc.F[b].z, Pk.R[1] * acc.F[b].x + Pk.R[4] * acc.F[b].y + Pk.R[7] * acc.F[b].z, Pk.R[2] * acc.F[b].x + Pk.R[5] * acc.F[b].y + Pk.R[8] * acc.F[b].z};
      const double* qh = A.qn.data() + B.qoff;
      V3d v{qh[1], qh[2], qh[3]}; double w = qh[0];
      V3d gv = scale(Gp, w) + cross(Gp, v);
      double inv = 2.0 / A.qnorm[b];
      grad[B.qoff] += -dot(Gp, v) * inv;
      grad[B.qoff + 1] += gv.x * inv; grad[B.qoff + 2] += gv.y * inv; grad[B.qoff + 3] += gv.z * inv;
      grad[B.qoff + 4] += Fp.x; grad[B.qoff + 5] += Fp.y; grad[B.qoff + 6] += Fp.z;
    }
  }
  double reg = 0;
  for (int k = m.nq; k < nx; ++k) { reg += x[k] * x[k]; grad[k] += 2.0 * w_def * x[k]; }
  acc.cost += w_def * reg;
}

// ---------------------------------------------------------------- depth rendering
// generateKinectRays (src/depthsensors.jl:10-30).  Note the reference multiplies x by tan(vertical_fov).
inline void kinect_rays(int rows, int cols, double vfov, double hfov, double* rays) {
  double cx = (cols + 1) / 2.0, cy = (rows + 1) / 2.0;
  double tv = std::tan(vfov), th = std::tan(hfov);
  for (int v = 1; v <= rows; ++v)
    for (int u = 1; u <= cols; ++u) {
      double r[3] = {(u - cx) * tv / cx, (v - cy) * th / cy, 1.0};
      double l = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
      double* o = rays + 3 * ((size_t)(v - 1) * cols + (u - 1));
      o[0] = r[0] / l; o[1] = r[1] / l; o[2] = r[2] / l;
    }
}

// doRaycast (src/depthsensors.jl:56-81)
template <class F> inline double do_raycast(const V3d& origin, const V3d& ray, F&& field) {
  const double EPS = 1e-5, SAFE_RATE = 0.4; const int SAFE_ITER_LIMIT = 60;
  double dist = 0; int k = 0; double est = -1;
  double last = field(origin + scale(ray, dist));
  while (std::fabs(last) > EPS && k < SAFE_ITER_LIMIT) {
    double step = -last / est;
    double sg = step > 0 ? 1.0 : (step < 0 ? -1.0 : 0.0);
    step = sg * std::min(SAFE_RATE, std::fabs(step));
    dist += step;
    double v = field(origin + scale(ray, dist));
    est = (v - last) / step;
    last = v;
    ++k;
  }
  if (std::fabs(field(origin + scale(ray, dist))) > 1000 * EPS) return std::numeric_limits<double>::quiet_NaN();
  return dist;
}

}  // namespace fo
