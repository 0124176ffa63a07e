// flash_host.h -- host-side model ingestion for libflashsdf: validates a flash_model_desc, builds
// the convex hulls (supporting planes) and their closest-feature BSP trees (flash_bsp.h), re-centres
// their geometry frames and packs everything into the device layout of flash_dev.h.  Plain C++17 (no CUDA), so the same code
// is exercised by the CPU-only tests.
//
// Replaces, for the hot path, what the Julia side holds in `Manipulator` / `ConvexGeometry` /
// `*InterpolatingSkin` (src/Flash.jl:35-65) and `collision_geometry` (src/models.jl:152).
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/flashsdf.h"
#include "flash_dev.h"
#include "flash_bsp.h"

namespace flashhost {

using namespace flashdev;

struct HostModel {
  ModelC M;
  std::vector<F4> geo_f4;
  std::vector<uint32_t> geo_u32;
  flash_options opt;
  std::vector<std::array<double, 5>> bsp_stats;   // per convex slot: nodes, max depth, mean leaf depth, cells, dropped slivers
};

struct HullOut {
  std::vector<D3> verts;                        // welded, re-centred
  D3 centroid;                                  // removed offset
  std::vector<std::array<double, 4>> planes;    // n.x - o <= 0 inside
};

inline double nrm(D3 a) { return std::sqrt(dot(a, a)); }

// Gift wrapping over vertex pairs: for every ordered pair (i, j) rotate a plane about the edge
// until no vertex lies in front of it; the pair is a hull edge iff the final plane supports the
// whole set.  O(n^3), n <= 255.
inline bool build_hull(const std::vector<D3>& in, HullOut& H, std::string& err) {
  double diam = 0;
  for (size_t i = 0; i < in.size(); ++i)
    for (size_t j = i + 1; j < in.size(); ++j) diam = std::max(diam, nrm(in[i] - in[j]));
  if (!(diam > 0)) { err = "convex surface has no extent"; return false; }
  std::vector<D3> v;
  for (auto& p : in) {
    bool dup = false;
    for (auto& q : v) if (nrm(p - q) <= 1e-9 * diam) { dup = true; break; }
    if (!dup) v.push_back(p);
  }
  const int n = (int)v.size();
  if (n < 4) { err = "convex surface needs at least 4 distinct vertices"; return false; }
  if (n > MAX_HULL_VERTS) { err = "convex surface has more than 255 distinct vertices"; return false; }
  D3 c = d3(0, 0, 0);
  for (auto& p : v) c = c + p;
  c = c * (1.0 / n);
  for (auto& p : v) p = p - c;
  H.centroid = c;
  const double eps = 1e-9 * diam;
  std::vector<std::array<double, 4>> planes;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      if (i == j) continue;
      D3 e = v[j] - v[i];
      D3 nn = d3(0, 0, 0); bool have = false;
      for (int m = 0; m < n; ++m) {
        if (m == i || m == j) continue;
        D3 r = v[m] - v[i];
        if (!have) {
          D3 t = cross(e, r);
          if (nrm(t) > 1e-12 * diam * diam) { nn = t * (1.0 / nrm(t)); have = true; }
          continue;
        }
        if (dot(nn, r) > eps) {
          D3 t = cross(e, r);
          double l = nrm(t);
          if (l > 1e-12 * diam * diam) nn = t * (1.0 / l);
        }
      }
      if (!have) continue;
      double o = dot(nn, v[i]);
      bool ok = true;
      for (int m = 0; m < n && ok; ++m) if (dot(nn, v[m]) - o > eps) ok = false;
      if (!ok) continue;
      bool seen = false;
      for (auto& p : planes)
        if (std::fabs(p[0] - nn.x) + std::fabs(p[1] - nn.y) + std::fabs(p[2] - nn.z) < 1e-7 && std::fabs(p[3] - o) < 1e-7 * diam) { seen = true; break; }
      if (!seen) planes.push_back({nn.x, nn.y, nn.z, o});
    }
  if (planes.size() < 4) { err = "convex surface is degenerate (flat)"; return false; }
  if (planes.size() > 1024) { err = "convex surface has more than 1024 faces"; return false; }
  H.verts = v;
  H.planes = planes;
  return true;
}

inline void options_default(flash_options* o) {
  std::memset(o, 0, sizeof *o);
  o->abi_version = FLASHSDF_ABI_VERSION;
  o->device = -1;
  o->rbf_normalize_by_gradient = 1;
  o->rbf_polynomial_degree = 1;
  o->sdf_mode = FLASH_SDF_EXACT;
  o->grid_spacing = 0.004;
  o->grid_margin = 0.10;
}

inline bool build_model(const flash_model_desc* d, const flash_options* opt_in, HostModel& out, std::string& err) {
  if (!d) { err = "null model description"; return false; }
  if (d->abi_version != FLASHSDF_ABI_VERSION) { err = "flash_model_desc.abi_version mismatch"; return false; }
  flash_options opt;
  if (opt_in) { if (opt_in->abi_version != FLASHSDF_ABI_VERSION) { err = "flash_options.abi_version mismatch"; return false; } opt = *opt_in; }
  else options_default(&opt);
  if (!(opt.grid_spacing > 0)) opt.grid_spacing = 0.004;
  if (!(opt.grid_margin > 0)) opt.grid_margin = 0.10;
  out.opt = opt;
  ModelC& M = out.M;
  std::memset(&M, 0, sizeof M);
  if (d->n_bodies < 0 || d->n_bodies > MAX_BODIES) { err = "too many bodies (max 32)"; return false; }
  if (d->n_surfaces < 1 || d->n_surfaces > MAX_SURFACES) { err = "need 1..32 surfaces"; return false; }
  if (d->n_bodies > 0 && (!d->body_parent || !d->joint_type || !d->joint_axis || !d->joint_to_parent)) { err = "null body arrays"; return false; }
  if (!d->surface_kind || !d->surface_body || !d->surface_frame || !d->point_offset || !d->n_surface_points || !d->points) { err = "null surface arrays"; return false; }
  M.n_bodies = d->n_bodies; M.n_surfaces = d->n_surfaces;
  M.normalize = opt.rbf_normalize_by_gradient ? 1 : 0;
  M.poly_deg = opt.rbf_polynomial_degree >= 1 ? 1 : 0;
  int qoff = 0;
  for (int b = 0; b < d->n_bodies; ++b) {
    BodyC& B = M.body[b];
    B.parent = d->body_parent[b];
    if (B.parent < -1 || B.parent >= b) { err = "bodies must be listed parents-first (body_parent[b] < b)"; return false; }
    B.jtype = d->joint_type[b];
    if (B.jtype < 0 || B.jtype > 2) { err = "unknown joint type"; return false; }
    B.qoff = qoff;
    qoff += B.jtype == JOINT_REVOLUTE ? 1 : B.jtype == JOINT_QUAT ? 7 : 0;
    double ax = d->joint_axis[3 * b], ay = d->joint_axis[3 * b + 1], az = d->joint_axis[3 * b + 2];
    double l = std::sqrt(ax * ax + ay * ay + az * az);
    if (B.jtype == JOINT_REVOLUTE && !(l > 0)) { err = "revolute joint with zero axis"; return false; }
    if (l > 0) { B.axis[0] = ax / l; B.axis[1] = ay / l; B.axis[2] = az / l; }
    std::memcpy(B.j2p, d->joint_to_parent + 12 * b, sizeof B.j2p);
  }
  if (qoff > 64) { err = "more than 64 configuration variables"; return false; }
  M.nq = qoff;
  int ndef = 0, tab = 0, acc = 1, auxs = 0, ptoff = 0, tabf = 0;
  // first pass: count convex surfaces so that the table offsets of skins follow them
  int ncv = 0;
  for (int s = 0; s < d->n_surfaces; ++s) if (d->surface_kind[s] == FLASH_SURF_CONVEX) ++ncv;
  if (ncv > MAX_CONVEX) { err = "too many convex surfaces (max 24)"; return false; }
  tab = 12 * ncv; acc = 1 + 6 * ncv; tabf = (12 * ncv + 3) & ~3;
  for (int s = 0; s < d->n_surfaces; ++s) {
    const int kind = d->surface_kind[s];
    const int p0 = d->point_offset[s], p1 = d->point_offset[s + 1];
    if (p1 < p0) { err = "point_offset must be non-decreasing"; return false; }
    M.surf_kind[s] = kind;
    if (kind == FLASH_SURF_CONVEX) {
      int j = M.n_convex++;
      M.surf_slot[s] = j;
      ConvexC& C = M.cvx[j];
      C.surface = s; C.body = d->surface_body[s];
      if (C.body < -1 || C.body >= d->n_bodies) { err = "convex surface attached to unknown body"; return false; }
      std::vector<D3> pts;
      for (int p = p0; p < p1; ++p) pts.push_back(d3(d->points[3 * p], d->points[3 * p + 1], d->points[3 * p + 2]));
      HullOut H;
      if (!build_hull(pts, H, err)) { err = "surface " + std::to_string(s) + ": " + err; return false; }
      C.nverts = (int)H.verts.size(); C.nplanes = (int)H.planes.size();
      for (int k = 0; k < 3; ++k) { C.lo[k] = 3e38f; C.hi[k] = -3e38f; }
      for (auto& v : H.verts) {
        const float f[3] = {(float)v.x, (float)v.y, (float)v.z};
        // bounds must contain the fp32 vertices (nextafter outward)
        for (int k = 0; k < 3; ++k) { C.lo[k] = std::min(C.lo[k], std::nextafterf(f[k], -3e38f)); C.hi[k] = std::max(C.hi[k], std::nextafterf(f[k], 3e38f)); }
      }
      {
        double rin = 1e300;   // the geometry origin is the vertex centroid: strictly inside
        for (auto& p : H.planes) rin = std::min(rin, p[3]);
        C.r_in = (float)std::max(0.0, rin * (1.0 - 1e-6) - 1e-7);
      }
      double db = 1e300;
      for (auto& p : H.planes) {
        double w = 0;
        for (auto& v : H.verts) w = std::max(w, p[3] - (p[0] * v.x + p[1] * v.y + p[2] * v.z));
        db = std::min(db, w);
      }
      C.depth_bound = (float)(db * 1.0001 + 1e-6);
      {
        // closest-feature BSP tree (flash_bsp.h) -> blob: node planes, plane table {n, -o}, vertices {x, y, z, 0}, edges; child words
        flashbsp::BspHull bh; std::string berr;
        if (!flashbsp::build_bsp(H.verts, H.planes, bh, berr)) { err = "surface " + std::to_string(s) + ": " + berr; return false; }
        C.n_nodes = (int)bh.node_plane.size();
        C.off_bn = (int)out.geo_f4.size();
        for (auto& p : bh.node_plane) out.geo_f4.push_back(F4{p[0], p[1], p[2], p[3]});
        C.off_bp = (int)out.geo_f4.size();
        for (auto& p : H.planes) out.geo_f4.push_back(F4{(float)p[0], (float)p[1], (float)p[2], -(float)p[3]});
        C.off_bv = (int)out.geo_f4.size();
        for (auto& v : H.verts) out.geo_f4.push_back(F4{(float)v.x, (float)v.y, (float)v.z, 0.f});
        C.off_be = (int)out.geo_f4.size();
        for (auto& e : bh.edges) out.geo_f4.push_back(F4{e[0], e[1], e[2], e[3]});
        C.off_bk = (int)out.geo_u32.size();
        for (uint32_t w : bh.node_kids) out.geo_u32.push_back(w);
        out.bsp_stats.push_back({(double)bh.node_plane.size(), (double)bh.max_depth, bh.mean_depth, (double)bh.n_regions, (double)bh.n_dropped});
      }
      // geometry frame = surface frame shifted to the centroid
      double fr[12], sh[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, H.centroid.x, H.centroid.y, H.centroid.z};
      xf_from_rowmajor34(d->surface_frame + 12 * s, fr);
      xf_compose(fr, sh, C.frame);
      C.g_on = 0;
    } else if (kind == FLASH_SURF_RIGID_SKIN || kind == FLASH_SURF_DEFORMABLE_SKIN) {
      if (M.n_skins >= MAX_SKINS) { err = "too many interpolating skins (max 4)"; return false; }
      int k = M.n_skins++;
      M.surf_slot[s] = k;
      SkinC& K = M.skin[k];
      K.surface = s; K.kind = kind; K.npts = p1 - p0; K.nsurf = d->n_surface_points[s];
      if (K.npts < 2 || K.npts > MAX_SKIN_PTS) { err = "skin needs 2..64 control points"; return false; }
      if (K.nsurf < 1 || K.nsurf > K.npts) { err = "skin n_surface_points out of range"; return false; }
      if (!d->point_body) { err = "skins need point_body"; return false; }
      K.M = K.npts + (M.poly_deg >= 1 ? 4 : 1);
      K.deform_off = -1;
      if (kind == FLASH_SURF_DEFORMABLE_SKIN) { K.deform_off = 3 * ndef; ndef += K.nsurf; }
      K.tab_off = tab; tab += 4 * K.npts + 4;
      K.acc_off = acc; acc += 4 * K.npts + 4;
      K.aux_off = auxs; auxs += K.M * K.M + K.M;
      K.pt_off = ptoff; ptoff += K.npts;
      K.tabf_off = tabf; tabf += 4 * K.npts + 4;
      for (int p = p0; p < p1; ++p) {
        int b = d->point_body[p];
        if (b < -1 || b >= d->n_bodies) { err = "skin control point on unknown body"; return false; }
        for (int c = 0; c < 3; ++c) M.skin_pt[K.pt_off + (p - p0)][c] = d->points[3 * p + c];
        M.skin_pt_body[K.pt_off + (p - p0)] = b;
      }
    } else { err = "unknown surface kind"; return false; }
  }
  M.nx = M.nq + 3 * ndef;
  M.tab_len = tab; M.acc_len = acc; M.aux_len = 25 * M.n_bodies + auxs; M.tabf_len = tabf;
  M.geo_f4 = (int)out.geo_f4.size(); M.geo_u32 = (int)out.geo_u32.size();
  return true;
}

inline void make_hot(const ModelC& M, HotModel& H) {
  std::memset(&H, 0, sizeof H);
  H.n_convex = M.n_convex; H.n_skins = M.n_skins; H.normalize = M.normalize; H.acc_len = M.acc_len; H.tab_len = M.tab_len; H.tabf_len = M.tabf_len;
  for (int k = 0; k < MAX_SKINS; ++k) H.skin[k] = M.skin[k];
  for (int j = 0; j < MAX_CONVEX; ++j) H.cvx[j] = M.cvx[j];
}

// generateKinectRays, src/depthsensors.jl:10-30 (the reference multiplies x by tan(vertical_fov))
inline void kinect_rays(int rows, int cols, double vfov, double hfov, double* rays) {
  const double cx = (cols + 1) / 2.0, cy = (rows + 1) / 2.0;
  const double tv = std::tan(vfov), th = std::tan(hfov);
  for (int v = 1; v <= rows; ++v)
    for (int u = 1; u <= cols; ++u) {
      double r0 = (u - cx) * tv / cx, r1 = (v - cy) * th / cy, r2 = 1.0;
      double l = std::sqrt(r0 * r0 + r1 * r1 + r2 * r2);
      double* o = rays + 3 * ((size_t)(v - 1) * cols + (u - 1));
      o[0] = r0 / l; o[1] = r1 / l; o[2] = r2 / l;
    }
}

}  // namespace flashhost
