// flash_urdf.h -- URDF + mesh ingestion behind the C ABI (plain C++17, no CUDA, no third-party parser).
//
// Replaces `load_urdf(filename; package_path)` of the reference (src/models.jl:154-171: RigidBodyDynamics.parse_urdf +
// RigidBodyTreeInspector.parse_urdf_visuals + extract_convex_surfaces) for callers that have no Julia: the kinematic
// tree in breadth-first order from the single root link (welded to the world, as parse_urdf does), one convex surface
// per <visual> mesh whose frame is the visual <origin> (the meshes of irb_140_convhull.urdf live in a shared CAD
// frame: dropping the origin silently breaks the model), vertices welded exactly (first occurrence kept),
// `package://` file names resolved against the given directories.  The result is a flash_model_desc that
// flash_model_create consumes.
#pragma once
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/flashsdf.h"

namespace flashurdf {

// ---------------------------------------------------------------------------------------------- a minimal XML reader
struct XmlNode {
  std::string name;
  std::map<std::string, std::string> attr;
  std::vector<std::unique_ptr<XmlNode>> kids;
  const XmlNode* child(const std::string& n) const { for (auto& k : kids) if (k->name == n) return k.get(); return nullptr; }
  std::vector<const XmlNode*> children(const std::string& n) const { std::vector<const XmlNode*> o; for (auto& k : kids) if (k->name == n) o.push_back(k.get()); return o; }
  const std::string* get(const std::string& a) const { auto it = attr.find(a); return it == attr.end() ? nullptr : &it->second; }
};

class XmlParser {
 public:
  explicit XmlParser(const std::string& text) : s(text) {}
  // returns the document's root element or null (err set)
  std::unique_ptr<XmlNode> parse(std::string& err) {
    std::unique_ptr<XmlNode> root;
    std::vector<XmlNode*> stack;
    while (true) {
      size_t lt = s.find('<', i);
      if (lt == std::string::npos) break;
      i = lt;
      if (s.compare(i, 4, "<!--") == 0) { size_t e = s.find("-->", i + 4); if (e == std::string::npos) { err = "unterminated comment"; return nullptr; } i = e + 3; continue; }
      if (s.compare(i, 2, "<?") == 0) { size_t e = s.find("?>", i + 2); if (e == std::string::npos) { err = "unterminated declaration"; return nullptr; } i = e + 2; continue; }
      if (s.compare(i, 9, "<![CDATA[") == 0) { size_t e = s.find("]]>", i + 9); if (e == std::string::npos) { err = "unterminated CDATA"; return nullptr; } i = e + 3; continue; }
      if (s.compare(i, 2, "<!") == 0) { size_t e = s.find('>', i + 2); if (e == std::string::npos) { err = "unterminated <!"; return nullptr; } i = e + 1; continue; }
      if (s.compare(i, 2, "</") == 0) {
        size_t e = s.find('>', i + 2);
        if (e == std::string::npos || stack.empty()) { err = "unbalanced closing tag"; return nullptr; }
        std::string n = trim(s.substr(i + 2, e - i - 2));
        if (n != stack.back()->name) { err = "closing tag </" + n + "> does not match <" + stack.back()->name + ">"; return nullptr; }
        stack.pop_back();
        i = e + 1;
        continue;
      }
      // element
      ++i;
      std::unique_ptr<XmlNode> node(new XmlNode);
      node->name = read_name();
      if (node->name.empty()) { err = "empty tag name"; return nullptr; }
      bool self_closed = false;
      while (true) {
        skip_ws();
        if (i >= s.size()) { err = "unterminated tag <" + node->name; return nullptr; }
        if (s[i] == '>') { ++i; break; }
        if (s[i] == '/' && i + 1 < s.size() && s[i + 1] == '>') { i += 2; self_closed = true; break; }
        std::string an = read_name();
        skip_ws();
        if (an.empty() || i >= s.size() || s[i] != '=') { err = "malformed attribute in <" + node->name + ">"; return nullptr; }
        ++i; skip_ws();
        if (i >= s.size() || (s[i] != '"' && s[i] != '\'')) { err = "unquoted attribute value in <" + node->name + ">"; return nullptr; }
        const char q = s[i++];
        size_t e = s.find(q, i);
        if (e == std::string::npos) { err = "unterminated attribute value"; return nullptr; }
        node->attr[an] = s.substr(i, e - i);
        i = e + 1;
      }
      XmlNode* raw = node.get();
      if (stack.empty()) { if (root) { err = "more than one root element"; return nullptr; } root = std::move(node); }
      else stack.back()->kids.push_back(std::move(node));
      if (!self_closed) stack.push_back(raw);
    }
    if (!stack.empty()) { err = "unclosed element <" + stack.back()->name + ">"; return nullptr; }
    if (!root) err = "no root element";
    return root;
  }

 private:
  const std::string& s;
  size_t i = 0;
  static std::string trim(const std::string& t) { size_t a = t.find_first_not_of(" \t\r\n"), b = t.find_last_not_of(" \t\r\n"); return a == std::string::npos ? "" : t.substr(a, b - a + 1); }
  void skip_ws() { while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\r' || s[i] == '\n')) ++i; }
  std::string read_name() {
    skip_ws();
    size_t b = i;
    while (i < s.size() && !(s[i] == ' ' || s[i] == '\t' || s[i] == '\r' || s[i] == '\n' || s[i] == '=' || s[i] == '>' || s[i] == '/')) ++i;
    return s.substr(b, i - b);
  }
};

inline bool read_file(const std::string& path, std::string& out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::ostringstream ss; ss << f.rdbuf(); out = ss.str();
  return true;
}
inline bool file_exists(const std::string& p) { std::ifstream f(p, std::ios::binary); return (bool)f; }

inline bool parse_doubles(const std::string& t, double* out, int n) {
  const char* c = t.c_str();
  for (int k = 0; k < n; ++k) { char* e = nullptr; out[k] = std::strtod(c, &e); if (e == c) return false; c = e; }
  return true;
}

// ---------------------------------------------------------------------------------------------- meshes
// Welded vertices, in order of first occurrence (exact equality: the triangles of an STL repeat their corners bit for bit).
inline void weld(std::vector<double>& v) {
  std::vector<double> out;
  std::map<std::array<double, 3>, int> seen;
  for (size_t i = 0; i + 2 < v.size(); i += 3) {
    std::array<double, 3> p = {v[i], v[i + 1], v[i + 2]};
    if (seen.emplace(p, 1).second) { out.push_back(p[0]); out.push_back(p[1]); out.push_back(p[2]); }
  }
  v.swap(out);
}
// binary (80-byte header, u32 count, 50 bytes per triangle: normal, 3 vertices as little-endian fp32, u16) or ASCII STL
inline bool read_stl(const std::string& path, std::vector<double>& verts, std::string& err) {
  std::string d;
  if (!read_file(path, d)) { err = "cannot open " + path; return false; }
  verts.clear();
  uint32_t ntri = 0;
  if (d.size() >= 84) std::memcpy(&ntri, d.data() + 80, 4);
  if (d.size() >= 84 && d.size() == 84 + 50 * (size_t)ntri) {
    for (uint32_t t = 0; t < ntri; ++t)
      for (int k = 0; k < 9; ++k) { float f; std::memcpy(&f, d.data() + 84 + 50 * (size_t)t + 12 + 4 * k, 4); verts.push_back((double)f); }
  } else {
    std::istringstream ss(d);
    std::string line;
    while (std::getline(ss, line)) {
      std::istringstream ls(line);
      std::string w; double x, y, z;
      if ((ls >> w) && w == "vertex" && (ls >> x >> y >> z)) { verts.push_back(x); verts.push_back(y); verts.push_back(z); }
    }
  }
  if (verts.empty()) { err = "no vertices in STL " + path; return false; }
  weld(verts);
  return true;
}
inline bool read_obj(const std::string& path, std::vector<double>& verts, std::string& err) {
  std::string d;
  if (!read_file(path, d)) { err = "cannot open " + path; return false; }
  verts.clear();
  std::istringstream ss(d);
  std::string line;
  while (std::getline(ss, line)) {
    if (line.size() < 2 || line[0] != 'v' || (line[1] != ' ' && line[1] != '\t')) continue;
    double p[3];
    if (!parse_doubles(line.substr(2), p, 3)) { err = "malformed vertex line in " + path; return false; }
    verts.insert(verts.end(), p, p + 3);
  }
  if (verts.empty()) { err = "no vertices in OBJ " + path; return false; }
  weld(verts);
  return true;
}

// ---------------------------------------------------------------------------------------------- URDF -> flash_model_desc
struct Urdf {
  std::vector<std::string> body_name;
  std::vector<int32_t> body_parent, joint_type;
  std::vector<double> joint_axis, joint_to_parent;
  std::vector<int32_t> surface_kind, surface_body, point_offset, n_surface_points, point_body;
  std::vector<double> surface_frame, points;
  std::vector<std::string> surface_mesh;    // resolved file of every surface
  flash_model_desc desc;
  void finish() {
    std::memset(&desc, 0, sizeof desc);
    desc.abi_version = FLASHSDF_ABI_VERSION;
    desc.n_bodies = (int32_t)body_parent.size();
    desc.body_parent = body_parent.data(); desc.joint_type = joint_type.data(); desc.joint_axis = joint_axis.data(); desc.joint_to_parent = joint_to_parent.data();
    desc.n_surfaces = (int32_t)surface_kind.size();
    desc.surface_kind = surface_kind.data(); desc.surface_body = surface_body.data(); desc.surface_frame = surface_frame.data();
    desc.point_offset = point_offset.data(); desc.n_surface_points = n_surface_points.data(); desc.points = points.data(); desc.point_body = point_body.data();
  }
};

// [R|t] row-major 3x4 of a URDF <origin xyz rpy>: R = Rz(yaw) Ry(pitch) Rx(roll)
inline bool origin_of(const XmlNode* elem, double T[12], std::string& err) {
  const double I[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
  std::memcpy(T, I, sizeof I);
  const XmlNode* o = elem ? elem->child("origin") : nullptr;
  if (!o) return true;
  double xyz[3] = {0, 0, 0}, rpy[3] = {0, 0, 0};
  if (const std::string* a = o->get("xyz")) if (!parse_doubles(*a, xyz, 3)) { err = "malformed origin xyz"; return false; }
  if (const std::string* a = o->get("rpy")) if (!parse_doubles(*a, rpy, 3)) { err = "malformed origin rpy"; return false; }
  const double cr = std::cos(rpy[0]), sr = std::sin(rpy[0]), cp = std::cos(rpy[1]), sp = std::sin(rpy[1]), cy = std::cos(rpy[2]), sy = std::sin(rpy[2]);
  // Rz Ry Rx, multiplied out in the order numpy evaluates (Rz @ Ry) @ Rx so that the doubles agree with the Python mirror
  const double A[9] = {cy * cp, -sy, cy * sp, sy * cp, cy, sy * sp, -sp, 0, cp};   // Rz Ry
  const double Rx[9] = {1, 0, 0, 0, cr, -sr, 0, sr, cr};
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) T[4 * r + c] = A[3 * r] * Rx[c] + A[3 * r + 1] * Rx[3 + c] + A[3 * r + 2] * Rx[6 + c];
  T[3] = xyz[0]; T[7] = xyz[1]; T[11] = xyz[2];
  return true;
}

inline bool resolve(const std::string& filename, const std::string& urdf_dir, const std::vector<std::string>& package_dirs, std::string& out, std::string& err) {
  const std::string pre = "package://";
  if (filename.compare(0, pre.size(), pre) == 0) {
    const std::string rel = filename.substr(pre.size());
    std::vector<std::string> roots = package_dirs;
    roots.push_back(urdf_dir);
    for (auto& root : roots) {
      std::string cand = root + "/" + rel;
      if (file_exists(cand)) { out = cand; return true; }
      // a package directory entry that IS the package: drop the package name
      size_t sl = rel.find('/');
      if (sl != std::string::npos) { cand = root + "/" + rel.substr(sl + 1); if (file_exists(cand)) { out = cand; return true; } }
    }
    err = "cannot resolve " + filename + " in the package directories";
    return false;
  }
  out = (!filename.empty() && filename[0] == '/') ? filename : urdf_dir + "/" + filename;
  if (!file_exists(out)) { err = "cannot open mesh " + out; return false; }
  return true;
}

inline bool load(const std::string& path, const std::vector<std::string>& package_dirs, Urdf& U, std::string& err) {
  std::string text;
  if (!read_file(path, text)) { err = "cannot open " + path; return false; }
  XmlParser xp(text);
  std::unique_ptr<XmlNode> root = xp.parse(err);
  if (!root) { err = path + ": " + err; return false; }
  if (root->name != "robot") { err = path + ": root element is <" + root->name + ">, expected <robot>"; return false; }
  std::string urdf_dir = ".";
  { size_t sl = path.find_last_of('/'); if (sl != std::string::npos) urdf_dir = path.substr(0, sl); }
  auto links = root->children("link");
  auto joints = root->children("joint");
  std::map<std::string, const XmlNode*> link_by_name;
  for (auto l : links) { const std::string* n = l->get("name"); if (!n) { err = "link without a name"; return false; } link_by_name[*n] = l; }
  std::map<std::string, int> is_child;
  for (auto j : joints) {
    const XmlNode* c = j->child("child"); const XmlNode* p = j->child("parent");
    if (!c || !p || !c->get("link") || !p->get("link")) { err = "joint without parent/child link"; return false; }
    if (!link_by_name.count(*c->get("link")) || !link_by_name.count(*p->get("link"))) { err = "joint refers to an unknown link"; return false; }
    is_child[*c->get("link")] = 1;
  }
  std::vector<std::string> roots;
  for (auto l : links) if (!is_child.count(*l->get("name"))) roots.push_back(*l->get("name"));   // document order
  if (roots.size() != 1) { err = "URDF must have exactly one root link, found " + std::to_string(roots.size()); return false; }
  // breadth-first from the root; the root link is welded to the world (RigidBodyDynamics.parse_urdf)
  std::map<std::string, int> index;
  auto push_body = [&](const std::string& name, int parent, int jt, const double axis[3], const double T[12]) {
    index[name] = (int)U.body_parent.size();
    U.body_name.push_back(name); U.body_parent.push_back(parent); U.joint_type.push_back(jt);
    U.joint_axis.insert(U.joint_axis.end(), axis, axis + 3);
    U.joint_to_parent.insert(U.joint_to_parent.end(), T, T + 12);
  };
  {
    const double ax[3] = {0, 0, 1}, I[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    push_body(roots[0], -1, FLASH_JOINT_FIXED, ax, I);
  }
  for (size_t q = 0; q < U.body_name.size(); ++q) {
    const std::string name = U.body_name[q];
    for (auto j : joints) {
      if (*j->child("parent")->get("link") != name) continue;
      const std::string child = *j->child("child")->get("link");
      if (index.count(child)) { err = "link " + child + " has more than one parent joint"; return false; }
      const std::string* jt = j->get("type");
      int kind;
      if (jt && (*jt == "revolute" || *jt == "continuous")) kind = FLASH_JOINT_REVOLUTE;
      else if (jt && *jt == "fixed") kind = FLASH_JOINT_FIXED;
      else if (jt && *jt == "floating") kind = FLASH_JOINT_QUAT_FLOATING;
      else { err = "unsupported URDF joint type '" + (jt ? *jt : std::string("")) + "'"; return false; }
      double axis[3] = {1, 0, 0};
      if (const XmlNode* a = j->child("axis")) if (const std::string* x = a->get("xyz")) if (!parse_doubles(*x, axis, 3)) { err = "malformed joint axis"; return false; }
      double T[12];
      if (!origin_of(j, T, err)) return false;
      push_body(child, (int)q, kind, axis, T);
    }
  }
  // one convex surface per <visual> mesh, links in mechanism order
  U.point_offset.push_back(0);
  for (size_t b = 0; b < U.body_name.size(); ++b) {
    for (auto vis : link_by_name[U.body_name[b]]->children("visual")) {
      const XmlNode* geo = vis->child("geometry");
      const XmlNode* mesh = geo ? geo->child("mesh") : nullptr;
      if (!mesh) continue;
      const std::string* fn = mesh->get("filename");
      if (!fn) { err = "mesh without a filename"; return false; }
      std::string file;
      if (!resolve(*fn, urdf_dir, package_dirs, file, err)) return false;
      std::vector<double> v;
      const bool obj = file.size() >= 4 && (file.compare(file.size() - 4, 4, ".obj") == 0 || file.compare(file.size() - 4, 4, ".OBJ") == 0);
      if (!(obj ? read_obj(file, v, err) : read_stl(file, v, err))) return false;
      if (const std::string* sc = mesh->get("scale")) {
        double k[3];
        if (!parse_doubles(*sc, k, 3)) { err = "malformed mesh scale"; return false; }
        for (size_t i = 0; i < v.size(); ++i) v[i] *= k[i % 3];
      }
      double T[12];
      if (!origin_of(vis, T, err)) return false;
      U.surface_kind.push_back(FLASH_SURF_CONVEX); U.surface_body.push_back((int32_t)b);
      U.surface_frame.insert(U.surface_frame.end(), T, T + 12);
      U.points.insert(U.points.end(), v.begin(), v.end());
      U.point_body.insert(U.point_body.end(), v.size() / 3, (int32_t)b);
      U.point_offset.push_back((int32_t)(U.points.size() / 3));
      U.n_surface_points.push_back((int32_t)(v.size() / 3));
      U.surface_mesh.push_back(file);
    }
  }
  if (U.surface_kind.empty()) { err = path + ": no <visual> mesh in any link"; return false; }
  U.finish();
  return true;
}

// ---------------------------------------------------------------------------------------------- point clouds
// read_point_cloud (src/depthdata.jl:19-30; examples/data/squishable_*_xyzrgb.txt): the first line is the camera
// origin `x, y, z`, every further line one return `x, y, z, r, g, b` (comma separated) -> origin[3], xyz[n][3], rgb[n][3].
inline bool read_point_cloud(const std::string& path, std::vector<double>& xyz, std::vector<double>* rgb, double origin[3], std::string& err) {
  std::ifstream f(path);
  if (!f) { err = "cannot open " + path; return false; }
  xyz.clear(); if (rgb) rgb->clear();
  std::string line;
  size_t ln = 0;
  bool have_origin = false;
  while (std::getline(f, line)) {
    ++ln;
    if (line.find_first_not_of(" \t\r") == std::string::npos) continue;
    for (auto& c : line) if (c == ',') c = ' ';
    double v[6];
    const char* c = line.c_str();
    int got = 0;
    for (; got < 6; ++got) { char* e = nullptr; v[got] = std::strtod(c, &e); if (e == c) break; c = e; }
    if (!have_origin) {
      if (got < 3) { err = path + ": the first line must hold the camera origin x, y, z"; return false; }
      if (origin) { origin[0] = v[0]; origin[1] = v[1]; origin[2] = v[2]; }
      have_origin = true;
      continue;
    }
    if (got < 6) { err = path + ": line " + std::to_string(ln) + " does not hold x, y, z, r, g, b"; return false; }
    xyz.insert(xyz.end(), v, v + 3);
    if (rgb) rgb->insert(rgb->end(), v + 3, v + 6);
  }
  if (!have_origin) { err = path + ": empty point cloud file"; return false; }
  return true;
}

}  // namespace flashurdf
