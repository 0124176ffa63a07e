// flash_bsp.h -- host-side builder of the closest-feature BSP tree of one convex hull (plain C++17, no CUDA).
//
// The signed distance to a convex polytope is a closed form of ONE feature inside every cell of a partition of
// space into convex cells:
//   outside   face f    : f's polygon swept along its normal                      d = n_f.y - o_f
//             edge (a,b): the edge plus the cone spanned by its two face normals   d = |(y - a) - ((y - a).u) u|
//             vertex v  : v plus its normal cone                                   d = |y - v|
//   inside    face f    : the points for which plane f is the nearest one          d = n_f.y - o_f  (< 0)
// (the reference computes the same number with EnhancedGJK, src/Flash.jl:238-243; inside, the exact depth is the
// stated deviation of DESIGN.md 2).  All cells are intersections of half-spaces, so a binary space partition whose
// cuts are cell-boundary planes locates the cell of a query point with one plane test per level: a branch-free,
// lane-uniform loop of ~10 instructions per level instead of GJK's support queries and simplex updates
// (flash_dev.h hull_sd_bsp; the per-point narrow phase was 45 % of cost_kernel's instructions at 17-23 of 32
// active lanes, profiles/r02_*).
//
// Construction: every cell is kept as an explicit convex polyhedron (faces = polygons, clipped to a large box);
// a tree node holds the cells that still overlap its region; the cut is the boundary plane that best balances the
// two sides (fewest cells on the fuller side, then fewest duplicated), cells are clipped by it, and parts thinner
// than `eps` (far below the fp32 resolution of the query) are dropped.  A query point within rounding distance of
// a cell boundary may be assigned to the neighbouring cell: the distance is continuous across every boundary (and
// C1 outside the hull), so the value changes by O(rounding) at most.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <cstdlib>
#include <string>
#include <vector>

#include "flash_dev.h"

namespace flashbsp {

using flashdev::D3;
using flashdev::d3;
using flashdev::dot;
using flashdev::cross;

inline double len(D3 a) { return std::sqrt(dot(a, a)); }

struct Plane { D3 n; double o; };   // unit n; "inside" of a half-space is n.x <= o

// convex polyhedron as a list of convex polygons; tag = id of the plane a face lies on (-1: bounding box)
struct Poly {
  std::vector<std::vector<D3>> faces;
  std::vector<int> tag;
  bool empty() const { return faces.size() < 4; }
};

inline Poly box_poly(double B) {
  Poly P;
  const D3 c[8] = {d3(-B, -B, -B), d3(B, -B, -B), d3(B, B, -B), d3(-B, B, -B), d3(-B, -B, B), d3(B, -B, B), d3(B, B, B), d3(-B, B, B)};
  const int f[6][4] = {{0, 3, 2, 1}, {4, 5, 6, 7}, {0, 1, 5, 4}, {2, 3, 7, 6}, {1, 2, 6, 5}, {0, 4, 7, 3}};
  for (auto& q : f) { P.faces.push_back({c[q[0]], c[q[1]], c[q[2]], c[q[3]]}); P.tag.push_back(-1); }
  return P;
}

// keep the part with s * (n.x - o) <= 0; `tag` names the cut face.  tol: vertices closer than this to the plane are on it.
inline void clip(Poly& P, const Plane& pl, double s, int tag, double tol) {
  if (P.empty()) { P.faces.clear(); P.tag.clear(); return; }
  std::vector<std::vector<D3>> nf; std::vector<int> nt;
  std::vector<D3> cut;
  bool any_out = false;
  for (size_t fi = 0; fi < P.faces.size(); ++fi) {
    const std::vector<D3>& F = P.faces[fi];
    const size_t m = F.size();
    std::vector<double> d(m);
    int n_in = 0, n_out = 0;
    for (size_t i = 0; i < m; ++i) { double v = s * (dot(pl.n, F[i]) - pl.o); if (std::fabs(v) <= tol) v = 0; d[i] = v; if (v < 0) ++n_in; else if (v > 0) ++n_out; }
    if (n_out == 0) {
      if (n_in == 0) { any_out = true; continue; }   // the face lies in the plane: regenerated as the cap
      nf.push_back(F); nt.push_back(P.tag[fi]);
      for (size_t i = 0; i < m; ++i) if (d[i] == 0) cut.push_back(F[i]);
      continue;
    }
    any_out = true;
    if (n_in == 0) { for (size_t i = 0; i < m; ++i) if (d[i] == 0) cut.push_back(F[i]); continue; }
    std::vector<D3> G;
    for (size_t i = 0; i < m; ++i) {
      const size_t j = (i + 1) % m;
      if (d[i] <= 0) G.push_back(F[i]);
      if (d[i] == 0) cut.push_back(F[i]);
      if ((d[i] < 0 && d[j] > 0) || (d[i] > 0 && d[j] < 0)) {
        const double t = d[i] / (d[i] - d[j]);
        const D3 x = F[i] + (F[j] - F[i]) * t;
        G.push_back(x); cut.push_back(x);
      }
    }
    if (G.size() >= 3) { nf.push_back(G); nt.push_back(P.tag[fi]); }
  }
  if (!any_out) return;   // entirely on the kept side
  // cap: the cut points, de-duplicated, ordered around their centroid in the plane
  std::vector<D3> u;
  for (auto& p : cut) { bool dup = false; for (auto& q : u) if (len(p - q) <= 4 * tol) { dup = true; break; } if (!dup) u.push_back(p); }
  if (u.size() >= 3) {
    D3 c = d3(0, 0, 0);
    for (auto& p : u) c = c + p;
    c = c * (1.0 / u.size());
    D3 e1 = d3(0, 0, 0);
    for (auto& p : u) if (len(p - c) > len(e1)) e1 = p - c;
    if (len(e1) > 0) {
      e1 = e1 * (1.0 / len(e1));
      const D3 nn = pl.n * s;                    // outward normal of the kept part at the cap
      const D3 e2 = cross(nn, e1);
      std::vector<std::pair<double, size_t>> ang;
      for (size_t i = 0; i < u.size(); ++i) ang.push_back({std::atan2(dot(u[i] - c, e2), dot(u[i] - c, e1)), i});
      std::sort(ang.begin(), ang.end());
      std::vector<D3> G;
      for (auto& a : ang) G.push_back(u[a.second]);
      // area check: a cap that is a segment is no face
      D3 ar = d3(0, 0, 0);
      for (size_t i = 1; i + 1 < G.size(); ++i) ar = ar + cross(G[i] - G[0], G[i + 1] - G[0]);
      if (len(ar) > tol * tol) { nf.push_back(G); nt.push_back(tag); }
    }
  }
  P.faces.swap(nf); P.tag.swap(nt);
  if (P.faces.size() < 4) { P.faces.clear(); P.tag.clear(); }
}

// volume / surface area: a lower bound of the inradius of a convex body (times 3: at least the inradius)
inline double thickness(const Poly& P) {
  if (P.empty()) return 0;
  D3 c = d3(0, 0, 0); size_t k = 0;
  for (auto& F : P.faces) for (auto& p : F) { c = c + p; ++k; }
  c = c * (1.0 / k);
  double V = 0, A = 0;
  for (auto& F : P.faces) {
    D3 ar = d3(0, 0, 0);
    for (size_t i = 1; i + 1 < F.size(); ++i) ar = ar + cross(F[i] - F[0], F[i + 1] - F[0]);
    A += 0.5 * len(ar);
    V += std::fabs(dot(ar, F[0] - c)) / 6.0;
  }
  return A > 0 ? 3.0 * V / A : 0.0;
}


// ---- hull topology: face polygons (counter-clockwise seen from outside) and edges with their two faces
struct Topology {
  std::vector<std::vector<int>> poly;          // per plane: vertex loop
  struct Edge { int a, b, f1, f2; };
  std::vector<Edge> edges;
  std::vector<std::vector<int>> vert_edges;    // per vertex: incident edges
};
inline bool build_topology(const std::vector<D3>& v, const std::vector<std::array<double, 4>>& planes, double diam, Topology& T, std::string& err) {
  const int np = (int)planes.size(), nv = (int)v.size();
  T.poly.assign(np, {}); T.edges.clear(); T.vert_edges.assign(nv, {});
  for (int f = 0; f < np; ++f) {
    const D3 n = d3(planes[f][0], planes[f][1], planes[f][2]);
    std::vector<int> on;
    for (int m = 0; m < nv; ++m) if (std::fabs(dot(n, v[m]) - planes[f][3]) <= 1e-7 * diam) on.push_back(m);
    if (on.size() < 3) { err = "bsp: face with fewer than 3 vertices"; return false; }
    // in-plane basis (e1, e2, n) right-handed: counter-clockwise in (e1, e2) is counter-clockwise seen from outside
    D3 e1 = std::fabs(n.x) < 0.9 ? cross(n, d3(1, 0, 0)) : cross(n, d3(0, 1, 0));
    e1 = e1 * (1.0 / len(e1));
    const D3 e2 = cross(n, e1);
    std::vector<std::array<double, 3>> p;   // x, y, id
    for (int m : on) p.push_back({dot(v[m], e1), dot(v[m], e2), (double)m});
    std::sort(p.begin(), p.end());
    auto turn = [](const std::array<double, 3>& a, const std::array<double, 3>& b, const std::array<double, 3>& c) { return (b[0] - a[0]) * (c[1] - a[1]) - (b[1] - a[1]) * (c[0] - a[0]); };
    const double ctol = 1e-9 * diam * diam;   // collinear (or interior) vertices are no polygon corners
    std::vector<std::array<double, 3>> h;
    for (int pass = 0; pass < 2; ++pass) {
      const size_t start = h.size();
      for (size_t i = 0; i < p.size(); ++i) {
        const auto& q = pass == 0 ? p[i] : p[p.size() - 1 - i];
        while (h.size() >= start + 2 && turn(h[h.size() - 2], h[h.size() - 1], q) <= ctol) h.pop_back();
        h.push_back(q);
      }
      h.pop_back();
    }
    if (h.size() < 3) { err = "bsp: degenerate face polygon"; return false; }
    for (auto& q : h) T.poly[f].push_back((int)q[2]);
  }
  for (int f = 0; f < np; ++f) {
    const std::vector<int>& L = T.poly[f];
    for (size_t i = 0; i < L.size(); ++i) {
      const int a = L[i], b = L[(i + 1) % L.size()];
      bool found = false;
      for (auto& e : T.edges) if (e.a == b && e.b == a) { if (e.f2 >= 0) { err = "bsp: edge shared by more than two faces"; return false; } e.f2 = f; found = true; break; }
      else if (e.a == a && e.b == b) { err = "bsp: inconsistent face orientation"; return false; }
      if (!found) T.edges.push_back({a, b, f, -1});
    }
  }
  for (size_t k = 0; k < T.edges.size(); ++k) {
    if (T.edges[k].f2 < 0) { err = "bsp: open edge (hull surface not closed)"; return false; }
    T.vert_edges[T.edges[k].a].push_back((int)k); T.vert_edges[T.edges[k].b].push_back((int)k);
  }
  return true;
}

// leaf codes: kind << 13 | index      kind 0: face (plane index), 1: vertex, 2: edge (index into BspHull::edges)
enum { LEAF_FACE = 0, LEAF_VERTEX = 1, LEAF_EDGE = 2 };
struct BspHull {
  std::vector<std::array<float, 4>> node_plane;   // {n, -o}: s = n.y + w;  s > 0 -> hi child, else lo child
  std::vector<uint32_t> node_kids;                // lo | hi << 16; bit 15 of a child: leaf, bits 0-14 its code
  std::vector<std::array<float, 4>> edges;        // {ux, uy, uz, bits(ia)}: unit direction, first vertex
  int max_depth = 0, n_leaves = 0, n_regions = 0, n_dropped = 0;
  double mean_depth = 0;                          // over leaves, unweighted
  double box = 0;                                 // half-width of the box the cells were built in (queries beyond it: see header)
};

struct Region { int code; std::vector<std::pair<int, double>> hs; };   // half-spaces: sign * (N[id].x - O[id]) <= 0

class Builder {
 public:
  std::vector<Plane> planes;     // global table of cut planes (de-duplicated up to orientation)
  std::vector<Region> regions;
  double scale = 1, tol = 1e-10, eps = 1e-8;
  std::pair<int, double> add_plane(D3 n, double o) {
    const double l = len(n);
    n = n * (1.0 / l); o /= l;
    for (size_t i = 0; i < planes.size(); ++i) {
      const Plane& p = planes[i];
      if (len(p.n - n) < 1e-9 && std::fabs(p.o - o) < 1e-9 * scale) return {(int)i, 1.0};
      if (len(p.n + n) < 1e-9 && std::fabs(p.o + o) < 1e-9 * scale) return {(int)i, -1.0};
    }
    planes.push_back({n, o});
    return {(int)planes.size() - 1, 1.0};
  }
  struct Item { int region; Poly poly; };
  std::vector<std::array<float, 4>> out_plane; std::vector<uint32_t> out_kids;
  int max_depth = 0, n_leaves = 0, n_dropped = 0; double depth_sum = 0;

  static void extent(const Poly& P, const Plane& pl, double& lo, double& hi) {
    lo = 1e300; hi = -1e300;
    for (auto& F : P.faces) for (auto& p : F) { const double d = dot(pl.n, p) - pl.o; lo = std::min(lo, d); hi = std::max(hi, d); }
  }
  uint32_t leaf(int code, int depth) { ++n_leaves; depth_sum += depth; max_depth = std::max(max_depth, depth); return 0x8000u | (uint32_t)code; }
  // returns the child word (leaf code | 0x8000, or node index)
  uint32_t build(std::vector<Item>& items, std::vector<char>& used, int depth) {
    bool same = true;
    for (auto& it : items) if (regions[it.region].code != regions[items[0].region].code) { same = false; break; }
    if (same || depth >= 60) {
      int best = 0; double bt = -1;
      if (!same) for (size_t i = 0; i < items.size(); ++i) { const double t = thickness(items[i].poly); if (t > bt) { bt = t; best = (int)i; } }
      return leaf(regions[items[best].region].code, depth);
    }
    // candidate cuts: planes that bound a cell part here and were not used on the path from the root
    std::vector<int> cand;
    {
      std::vector<char> seen(planes.size(), 0);
      for (auto& it : items) for (int t : it.poly.tag) if (t >= 0 && !used[t] && !seen[t]) { seen[t] = 1; cand.push_back(t); }
    }
    const size_t n = items.size();
    // near the root there are ~1000 candidate planes for ~400 cells: score an evenly strided sample of them
    if (cand.size() > 96) { std::vector<int> sub; const double st = (double)cand.size() / 96.0; for (int q = 0; q < 96; ++q) sub.push_back(cand[(size_t)(q * st)]); cand.swap(sub); }
    // the cells' distinct vertices (a polygon list names every vertex three times or more)
    std::vector<std::vector<D3>> pv(n);
    for (size_t i = 0; i < n; ++i) {
      for (auto& F : items[i].poly.faces) for (auto& p : F) pv[i].push_back(p);
      std::sort(pv[i].begin(), pv[i].end(), [](const D3& a, const D3& b) { return a.x != b.x ? a.x < b.x : (a.y != b.y ? a.y < b.y : a.z < b.z); });
      pv[i].erase(std::unique(pv[i].begin(), pv[i].end(), [](const D3& a, const D3& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }), pv[i].end());
    }
    int best = -1; long long best_score = -1;
    std::vector<double> lo(n), hi(n);
    const double side = 0.05 * eps;
    for (int c : cand) {
      long long nl = 0, nr = 0;
      for (size_t i = 0; i < n; ++i) {
        double a = 1e300, b = -1e300;
        for (auto& q : pv[i]) { const double dq = dot(planes[c].n, q) - planes[c].o; a = std::min(a, dq); b = std::max(b, dq); }
        if (a < -side) ++nl;
        if (b > side) ++nr;
        if (!(a < -side) && !(b > side)) { ++nl; }
      }
      if (nl >= (long long)n && nr >= (long long)n) continue;   // separates nothing
      static const double w_split = getenv("BSP_WSPLIT") ? atof(getenv("BSP_WSPLIT")) : 1.5;
      const long long score = (long long)(1000.0 * (w_split * (double)(nl + nr - (long long)n) + (double)std::llabs(nl - nr)));
      if (best < 0 || score < best_score) { best = c; best_score = score; }
    }
    if (best < 0) {   // cannot be separated (numerically coincident cells): the thickest part wins
      int bi = 0; double bt = -1;
      for (size_t i = 0; i < n; ++i) { const double t = thickness(items[i].poly); if (t > bt) { bt = t; bi = (int)i; } }
      n_dropped += (int)n - 1;
      return leaf(regions[items[bi].region].code, depth);
    }
    const Plane& pl = planes[best];
    std::vector<Item> L, R;
    for (auto& it : items) {
      double own = 0;
      for (auto& h : regions[it.region].hs) if (h.first == best) own = h.second;
      if (own > 0) { L.push_back(std::move(it)); continue; }
      if (own < 0) { R.push_back(std::move(it)); continue; }
      double a, b; extent(it.poly, pl, a, b);
      if (!(b > side)) { L.push_back(std::move(it)); continue; }
      if (!(a < -side)) { R.push_back(std::move(it)); continue; }
      Item l{it.region, it.poly}, r{it.region, std::move(it.poly)};
      clip(l.poly, pl, 1.0, best, tol); clip(r.poly, pl, -1.0, best, tol);
      const bool kl = thickness(l.poly) > eps, kr = thickness(r.poly) > eps;
      if (kl) L.push_back(std::move(l));
      if (kr) R.push_back(std::move(r));
      if (!kl && !kr) { ++n_dropped; }
    }
    items.clear(); items.shrink_to_fit();
    const uint32_t me = (uint32_t)out_plane.size();
    out_plane.push_back({(float)pl.n.x, (float)pl.n.y, (float)pl.n.z, (float)(-pl.o)});
    out_kids.push_back(0);
    used[best] = 1;
    uint32_t kl, kr;
    // an empty side can only be reached by points within rounding distance of the cut: send them to the other side's answer
    if (L.empty() && R.empty()) { used[best] = 0; return leaf(0, depth); }
    if (L.empty()) { kr = build(R, used, depth + 1); kl = kr; }
    else if (R.empty()) { kl = build(L, used, depth + 1); kr = kl; }
    else { kl = build(L, used, depth + 1); kr = build(R, used, depth + 1); }
    used[best] = 0;
    out_kids[me] = kl | (kr << 16);
    return me;
  }
};

inline bool build_bsp_uncached(const std::vector<D3>& v, const std::vector<std::array<double, 4>>& hp, BspHull& out, std::string& err);
// A process builds the same hulls again and again (one ctx per model, tests, multi-GPU ranks): the trees are cached by
// the exact bytes of their input.
inline bool build_bsp(const std::vector<D3>& v, const std::vector<std::array<double, 4>>& hp, BspHull& out, std::string& err) {
  static std::mutex mu;
  static std::map<std::string, BspHull> cache;
  std::string key((const char*)v.data(), v.size() * sizeof(D3));
  key.append((const char*)hp.data(), hp.size() * sizeof(hp[0]));
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { out = it->second; return true; }
  }
  if (!build_bsp_uncached(v, hp, out, err)) return false;
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() < 256) cache[key] = out;
  return true;
}
inline bool build_bsp_uncached(const std::vector<D3>& v, const std::vector<std::array<double, 4>>& hp, BspHull& out, std::string& err) {
  double diam = 0;
  for (size_t i = 0; i < v.size(); ++i) for (size_t j = i + 1; j < v.size(); ++j) diam = std::max(diam, len(v[i] - v[j]));
  Topology T;
  if (!build_topology(v, hp, diam, T, err)) return false;
  Builder B;
  B.scale = diam; B.tol = 1e-10 * std::max(diam, 1.0); B.eps = 2e-8 * diam;
  const double box = std::max(64.0, 256.0 * diam);
  out.box = box;
  const int np = (int)hp.size();
  auto P = [&](int f) { return d3(hp[f][0], hp[f][1], hp[f][2]); };
  std::vector<Builder::Item> items;
  auto add_region = [&](int code, std::vector<std::pair<int, double>> hs, const Poly* start) {
    Builder::Item it; it.region = (int)B.regions.size();
    it.poly = start ? *start : box_poly(box);
    for (auto& h : hs) clip(it.poly, B.planes[h.first], h.second, h.first, B.tol);
    B.regions.push_back({code, std::move(hs)});
    if (thickness(it.poly) > B.eps) items.push_back(std::move(it)); else ++out.n_dropped;
  };
  // face planes first: plane id f = hull plane f
  for (int f = 0; f < np; ++f) B.add_plane(P(f), hp[f][3]);
  // outside cells of faces, edges, vertices
  std::vector<std::vector<std::pair<int, double>>> side(np);   // per face: its prism side planes, oriented outwards
  for (int f = 0; f < np; ++f) {
    const std::vector<int>& L = T.poly[f];
    std::vector<std::pair<int, double>> hs;
    hs.push_back({f, -1.0});                                    // outside plane f
    for (size_t i = 0; i < L.size(); ++i) {
      const D3 a = v[L[i]], b = v[L[(i + 1) % L.size()]];
      D3 s = cross(b - a, P(f));                                // in-plane outward normal of edge a -> b
      s = s * (1.0 / len(s));
      const auto id = B.add_plane(s, dot(s, a));
      hs.push_back(id); side[f].push_back(id);
    }
    add_region((LEAF_FACE << 13) | f, hs, nullptr);
  }
  if (T.edges.size() >= (1u << 13) || np >= (1 << 13)) { err = "bsp: too many features"; return false; }
  for (size_t k = 0; k < T.edges.size(); ++k) {
    const auto& e = T.edges[k];
    const D3 a = v[e.a], b = v[e.b];
    D3 u = b - a; const double l = len(u); u = u * (1.0 / l);
    std::vector<std::pair<int, double>> hs;
    for (int f : {e.f1, e.f2}) {
      D3 s = (f == e.f1) ? cross(b - a, P(f)) : cross(a - b, P(f));   // the edge runs a -> b in f1 and b -> a in f2
      s = s * (1.0 / len(s));
      auto id = B.add_plane(s, dot(s, a));
      hs.push_back({id.first, -id.second});                      // beyond the prism side plane
    }
    auto lo_ = B.add_plane(u * -1.0, -dot(u, a)); hs.push_back(lo_);   // u.x >= u.a
    auto hi_ = B.add_plane(u, dot(u, b)); hs.push_back(hi_);          // u.x <= u.b
    add_region((LEAF_EDGE << 13) | (int)k, hs, nullptr);
    uint32_t ia = (uint32_t)e.a; float fa; std::memcpy(&fa, &ia, 4);
    out.edges.push_back({(float)u.x, (float)u.y, (float)u.z, fa});
  }
  for (size_t m = 0; m < v.size(); ++m) {
    if (T.vert_edges[m].empty()) continue;                      // not a corner of the hull
    std::vector<std::pair<int, double>> hs;
    for (int k : T.vert_edges[m]) {
      const auto& e = T.edges[k];
      const int w = e.a == (int)m ? e.b : e.a;
      const D3 u = v[w] - v[m];
      hs.push_back(B.add_plane(u, dot(u, v[m])));                // (w - v).(x - v) <= 0
    }
    add_region((LEAF_VERTEX << 13) | (int)m, hs, nullptr);
  }
  // inside cells: the hull clipped by the bisectors against every other face
  Poly hull = box_poly(box);
  for (int f = 0; f < np; ++f) clip(hull, B.planes[f], 1.0, f, B.tol);
  if (hull.empty()) { err = "bsp: empty hull polyhedron"; return false; }
  for (int f = 0; f < np; ++f) {
    std::vector<std::pair<int, double>> hs;
    Poly cell = hull;
    for (int g = 0; g < np && !cell.empty(); ++g) {
      if (g == f) continue;
      const D3 dn = P(g) - P(f);
      if (len(dn) < 1e-12) continue;
      // skip bisectors that do not touch what is left of the cell
      Plane pl{dn * (1.0 / len(dn)), (hp[g][3] - hp[f][3]) / len(dn)};
      double a, b; Builder::extent(cell, pl, a, b);
      if (!(b > 0)) continue;
      const auto id = B.add_plane(dn, hp[g][3] - hp[f][3]);
      clip(cell, B.planes[id.first], id.second, id.first, B.tol);
      hs.push_back(id);
    }
    hs.push_back({f, 1.0});
    Builder::Item it; it.region = (int)B.regions.size(); it.poly = std::move(cell);
    B.regions.push_back({(LEAF_FACE << 13) | f, std::move(hs)});
    if (thickness(it.poly) > B.eps) items.push_back(std::move(it)); else ++out.n_dropped;
  }
  out.n_regions = (int)items.size();
  std::vector<char> used(B.planes.size(), 0);
  used.reserve(B.planes.size());
  const uint32_t root = B.build(items, used, 0);
  if (root & 0x8000u) { err = "bsp: degenerate tree"; return false; }
  if (B.out_plane.size() >= 0x8000u) { err = "bsp: more than 32767 nodes"; return false; }
  out.node_plane.swap(B.out_plane); out.node_kids.swap(B.out_kids);
  out.max_depth = B.max_depth; out.n_leaves = B.n_leaves; out.n_dropped += B.n_dropped;
  out.mean_depth = B.n_leaves ? B.depth_sum / B.n_leaves : 0;
  return true;
}

}  // namespace flashbsp
