/* abi_driver.c -- TEST INFRASTRUCTURE: calls every entry point of include/flashsdf.h from plain C, the caller a Julia
 * `ccall` is closest to (no Python, no ctypes marshalling in between).  Built by __graft_entry__.build() with gcc and
 * run by tests/test_abi.py: without a GPU it checks the host-side entry points and that every compute call fails
 * loudly with FLASH_ERR_NO_DEVICE; with one it checks the results against closed forms (a box, a beanbag skin).
 *
 *   abi_driver <path to irb_140_convhull.urdf> <package dir> <scratch dir>
 * prints one line per check and exits 0 only if all passed. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/flashsdf.h"

static int failures = 0;
#define CHECK(cond, ...) do { if (cond) { printf("ok   "); } else { printf("FAIL "); ++failures; } printf(__VA_ARGS__); printf("\n"); } while (0)

/* a box (half extents 0.25, 0.25, 0.05: the table of examples/irb_and_squishable.ipynb) on a quaternion-floating body,
 * plus the beanbag skin of src/models.jl:85-93 on a second floating body */
static const double box_pts[8][3] = {{-0.25, -0.25, -0.05}, {-0.25, 0.25, -0.05}, {0.25, -0.25, -0.05}, {0.25, 0.25, -0.05},
                                     {-0.25, -0.25, 0.05},  {-0.25, 0.25, 0.05},  {0.25, -0.25, 0.05},  {0.25, 0.25, 0.05}};
static const double bag_pts[7][3] = {{-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}, {0, 0, 0}};

static double box_sd(const double p[3], const double c[3]) {
  const double h[3] = {0.25, 0.25, 0.05};
  double q[3], o2 = 0, m = -1e300;
  for (int k = 0; k < 3; ++k) { q[k] = fabs(p[k] - c[k]) - h[k]; if (q[k] > 0) o2 += q[k] * q[k]; if (q[k] > m) m = q[k]; }
  return sqrt(o2) + (m < 0 ? m : 0);
}

int main(int argc, char** argv) {
  const char* urdf = argc > 1 ? argv[1] : NULL;
  const char* pkg = argc > 2 ? argv[2] : NULL;
  const char* scratch = argc > 3 ? argv[3] : "/tmp";
  CHECK(flash_abi_version() == FLASHSDF_ABI_VERSION, "flash_abi_version() == %d", FLASHSDF_ABI_VERSION);
  flash_options opt;
  flash_options_default(&opt);
  CHECK(opt.abi_version == FLASHSDF_ABI_VERSION && opt.device == -1 && opt.rbf_normalize_by_gradient == 1 && opt.rbf_polynomial_degree == 1 &&
        opt.sdf_mode == FLASH_SDF_EXACT && opt.deterministic == 0 && opt.n_devices == 0, "flash_options_default");
  double rays[2 * 3 * 3];
  CHECK(flash_kinect_rays(2, 3, 0.4682, 0.5449, rays) == FLASH_OK && fabs(rays[0] * rays[0] + rays[1] * rays[1] + rays[2] * rays[2] - 1.0) < 1e-12, "flash_kinect_rays: unit rays");

  /* ---- model description by hand */
  int32_t parent[2] = {-1, -1}, jtype[2] = {FLASH_JOINT_QUAT_FLOATING, FLASH_JOINT_QUAT_FLOATING};
  double axis[6] = {0, 0, 1, 0, 0, 1};
  double j2p[24] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
  int32_t kind[2] = {FLASH_SURF_CONVEX, FLASH_SURF_RIGID_SKIN}, sbody[2] = {0, -1}, offs[3] = {0, 8, 15}, nsurf[2] = {8, 6};
  double sframe[24] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
  double pts[15][3];
  int32_t pbody[15];
  memcpy(pts, box_pts, sizeof box_pts); memcpy(pts + 8, bag_pts, sizeof bag_pts);
  for (int i = 0; i < 15; ++i) pbody[i] = i < 8 ? 0 : 1;
  flash_model_desc d;
  memset(&d, 0, sizeof d);
  d.abi_version = FLASHSDF_ABI_VERSION; d.n_bodies = 2; d.body_parent = parent; d.joint_type = jtype; d.joint_axis = axis; d.joint_to_parent = j2p;
  d.n_surfaces = 2; d.surface_kind = kind; d.surface_body = sbody; d.surface_frame = sframe; d.point_offset = offs; d.n_surface_points = nsurf;
  d.points = &pts[0][0]; d.point_body = pbody;

  /* ---- files: URDF + mesh ingestion and the point-cloud reader need no device */
  if (urdf) {
    flash_urdf* u = NULL;
    const char* dirs[1] = {pkg};
    int rc = flash_urdf_load(urdf, dirs, pkg ? 1 : 0, &u);
    CHECK(rc == FLASH_OK && u, "flash_urdf_load(%s): %s", urdf, rc ? flash_last_error(NULL) : "ok");
    if (u) {
      const flash_model_desc* ud = flash_urdf_desc(u);
      CHECK(ud->n_bodies == 7 && ud->n_surfaces == 7 && ud->point_offset[7] == 7 * 52, "IRB140: 7 bodies, 7 surfaces, 52 welded vertices each");
      CHECK(strcmp(flash_urdf_body_name(u, 0), "base_link") == 0 && strcmp(flash_urdf_body_name(u, 6), "link_6") == 0 && flash_urdf_body_name(u, 7) == NULL, "flash_urdf_body_name");
      CHECK(strstr(flash_urdf_surface_mesh(u, 1), "link1_chull.stl") != NULL, "flash_urdf_surface_mesh");
      CHECK(flash_urdf_set_joint_type(u, 0, FLASH_JOINT_QUAT_FLOATING) == FLASH_OK && ud->joint_type[0] == FLASH_JOINT_QUAT_FLOATING &&
            flash_urdf_set_joint_type(u, 0, 9) == FLASH_ERR_INVALID, "flash_urdf_set_joint_type");
      flash_urdf_free(u);
    }
    CHECK(flash_urdf_load("/nonexistent.urdf", NULL, 0, &u) == FLASH_ERR_IO && u == NULL && strstr(flash_last_error(NULL), "cannot open"), "flash_urdf_load: missing file -> FLASH_ERR_IO");
  }
  char cloud_path[1024];
  snprintf(cloud_path, sizeof cloud_path, "%s/abi_driver_cloud.txt", scratch);
  {
    FILE* f = fopen(cloud_path, "w");
    fprintf(f, "0.5, 0.25, 2.0\n");
    for (int i = 0; i < 300; ++i) fprintf(f, "%.17g, %.17g, %.17g, 0.1, 0.2, 0.3\n", 0.004 * i - 0.6, 0.3 * sin(0.1 * i), 0.2 + 0.002 * i);
    fclose(f);
    double origin[3], xyz[300 * 3], rgb[300 * 3]; int64_t n = 0;
    CHECK(flash_read_point_cloud(cloud_path, origin, NULL, NULL, 0, &n) == FLASH_OK && n == 300, "flash_read_point_cloud: count-only call");
    CHECK(flash_read_point_cloud(cloud_path, origin, xyz, rgb, 300, &n) == FLASH_OK && origin[2] == 2.0 && xyz[3 * 299] == 0.004 * 299 - 0.6 && rgb[5] == 0.3, "flash_read_point_cloud: values");
    CHECK(flash_read_point_cloud("/nonexistent.txt", origin, NULL, NULL, 0, &n) == FLASH_ERR_IO, "flash_read_point_cloud: missing file -> FLASH_ERR_IO");
  }

  /* ---- context */
  flash_ctx* ctx = NULL;
  int rc = flash_model_create(&d, &opt, &ctx);
  if (rc == FLASH_ERR_NO_DEVICE) {
    CHECK(ctx == NULL && strstr(flash_last_error(NULL), "no CPU path") != NULL, "no CUDA device: flash_model_create fails loudly (%s)", flash_last_error(NULL));
    if (urdf) { const char* dirs[1] = {pkg}; CHECK(flash_model_load_urdf(urdf, dirs, pkg ? 1 : 0, &opt, &ctx) == FLASH_ERR_NO_DEVICE, "flash_model_load_urdf: FLASH_ERR_NO_DEVICE"); }
    d.abi_version = 77;
    CHECK(flash_model_create(&d, &opt, &ctx) == FLASH_ERR_INVALID, "bad abi_version rejected before the device is touched");
    CHECK(flash_num_states(NULL) == FLASH_ERR_INVALID && flash_cost_grad(NULL, NULL, 1, 0, NULL, NULL, NULL) == FLASH_ERR_INVALID && flash_points_swap(NULL) == FLASH_ERR_INVALID, "null ctx -> FLASH_ERR_INVALID");
    flash_destroy(NULL);
    printf("%s (host-side checks only: no CUDA device)\n", failures ? "FAILED" : "PASSED");
    return failures ? 1 : 0;
  }
  CHECK(rc == FLASH_OK && ctx, "flash_model_create: %s", rc ? flash_last_error(NULL) : "ok");
  if (!ctx) return 1;
  CHECK(flash_num_positions(ctx) == 14 && flash_num_states(ctx) == 14 && flash_num_surfaces(ctx) == 2 && flash_num_devices(ctx) == 1 && flash_comm_world(ctx) == 1, "counts");
  int32_t sms = 0, khz = 0; int64_t l2 = 0;
  CHECK(flash_device_info(ctx, &sms, &khz, &l2) == FLASH_OK && sms > 0 && l2 > 0, "flash_device_info: %d SMs, L2 %lld MB", sms, (long long)(l2 >> 20));
  double x[14] = {1, 0, 0, 0, 0.1, -0.2, 0.3, 1, 0, 0, 0, 5, 5, 5};   /* box at (0.1,-0.2,0.3), bag far away */
  enum { N = 4096 };
  static double cloud[N][3];
  double ref = 0;
  const double c[3] = {0.1, -0.2, 0.3};
  for (int i = 0; i < N; ++i) {
    cloud[i][0] = (float)(c[0] + 0.8 * sin(0.37 * i)); cloud[i][1] = (float)(c[1] + 0.8 * sin(0.11 * i + 1)); cloud[i][2] = (float)(c[2] + 0.4 * sin(0.23 * i + 2));
    const double s = box_sd(cloud[i], c); ref += s * s;
  }
  double cost = 0, grad[14];
  CHECK(flash_cost_grad(ctx, x, 1, 10.0, &cost, grad, NULL) == FLASH_ERR_STATE, "cost before points -> FLASH_ERR_STATE (%s)", flash_last_error(ctx));
  CHECK(flash_points_set(ctx, &cloud[0][0], N) == FLASH_OK && flash_num_points(ctx) == N && flash_num_frames(ctx) == 1, "flash_points_set");
  static int32_t arg[N];
  CHECK(flash_profiling(ctx, 1) == FLASH_OK && flash_cost_grad(ctx, x, 1, 10.0, &cost, grad, arg) == FLASH_OK, "flash_cost_grad");
  CHECK(fabs(cost - ref) < 1e-5 * ref, "cost %.9g against the closed form of the box %.9g", cost, ref);
  int all0 = 1; for (int i = 0; i < N; ++i) all0 &= arg[i] == 0;
  CHECK(all0, "argmin: the box everywhere (the bag is 8 m away)");
  { /* gradient w.r.t. the box translation by central differences of the closed form */
    double fd[3];
    for (int k = 0; k < 3; ++k) {
      double cp[3] = {c[0], c[1], c[2]}, cm[3] = {c[0], c[1], c[2]}, fp = 0, fm = 0;
      cp[k] += 1e-6; cm[k] -= 1e-6;
      for (int i = 0; i < N; ++i) { double a = box_sd(cloud[i], cp), b = box_sd(cloud[i], cm); fp += a * a; fm += b * b; }
      fd[k] = (fp - fm) / 2e-6;
    }
    CHECK(fabs(grad[4] - fd[0]) + fabs(grad[5] - fd[1]) + fabs(grad[6] - fd[2]) < 1e-4 * (fabs(fd[0]) + fabs(fd[1]) + fabs(fd[2])), "d cost / d translation (%.6g %.6g %.6g) against central differences", grad[4], grad[5], grad[6]);
  }
  flash_timing tm;
  CHECK(flash_last_timing(ctx, &tm) == FLASH_OK && tm.ms_total > 0 && tm.launches >= 3 && flash_launch_count(ctx) >= tm.launches, "flash_last_timing: %.3f ms, %d launches", tm.ms_total, tm.launches);
  { /* skin_eval: beanbag known answer (test/runtests.jl:17: skin(100,0,0) ~ 99 within 2 %) with the bag at the origin */
    double xb[14] = {1, 0, 0, 0, 500, 500, 500, 1, 0, 0, 0, 0, 0, 0}, q[3] = {100, 0, 0}, dist = 0; int32_t a = -1;
    CHECK(flash_skin_eval(ctx, xb, q, 1, &dist, &a) == FLASH_OK && fabs(dist - 99.0) < 0.02 * 99.0 && a == 1, "flash_skin_eval: beanbag skin(100,0,0) = %.4f (reference test: 99 +- 2 %%)", dist);
    double d2 = 0; int32_t a2 = -1;
    CHECK(flash_skin_eval_static(ctx, xb, q, 1, &d2, &a2) == FLASH_OK && d2 == dist && a2 == a, "flash_skin_eval_static (no allocation): the same value, %.4f", d2);
    CHECK(flash_skin_eval_static(ctx, xb, q, FLASH_STATIC_MAX_QUERY + 1, &d2, &a2) == FLASH_ERR_TOO_LARGE && strstr(flash_last_error(ctx), "FLASH_STATIC_MAX_QUERY"), "flash_skin_eval_static: beyond its maximum -> FLASH_ERR_TOO_LARGE");
  }
  { /* float, file, async + swap entry points give the same cost (to rounding: the default mode sums in a different order every run) */
    static float cf[N][3];
    for (int i = 0; i < N; ++i) for (int k = 0; k < 3; ++k) cf[i][k] = (float)cloud[i][k];
    double c2 = 0, c3 = 0, c4 = 0, c5 = 0;
    CHECK(flash_points_set_f32(ctx, &cf[0][0], N) == FLASH_OK && flash_cost_grad(ctx, x, 1, 10.0, &c2, NULL, NULL) == FLASH_OK && fabs(c2 - cost) < 1e-12 * cost, "flash_points_set_f32");
    CHECK(flash_points_set_async(ctx, &cloud[0][0], N / 2) == FLASH_OK && flash_points_swap(ctx) == FLASH_OK && flash_points_sync(ctx) == FLASH_OK && flash_num_points(ctx) == N / 2 &&
          flash_points_swap(ctx) == FLASH_ERR_STATE, "flash_points_set_async / swap / sync");
    CHECK(flash_points_set_async_f32(ctx, &cf[0][0], N) == FLASH_OK && flash_points_swap(ctx) == FLASH_OK && flash_cost_grad(ctx, x, 1, 10.0, &c3, NULL, NULL) == FLASH_OK && fabs(c3 - cost) < 1e-12 * cost, "flash_points_set_async_f32");
    FILE* f = fopen(cloud_path, "w");
    fprintf(f, "0, 0, 0\n");
    for (int i = 0; i < N; ++i) fprintf(f, "%.9g, %.9g, %.9g, 0, 0, 0\n", cloud[i][0], cloud[i][1], cloud[i][2]);
    fclose(f);
    CHECK(flash_points_set_file(ctx, cloud_path, NULL) == FLASH_OK && flash_cost_grad(ctx, x, 1, 10.0, &c4, NULL, NULL) == FLASH_OK && fabs(c4 - cost) < 1e-12 * cost, "flash_points_set_file");
    CHECK(flash_points_set_file_async(ctx, cloud_path, NULL) == FLASH_OK && flash_points_swap(ctx) == FLASH_OK && flash_cost_grad(ctx, x, 1, 10.0, &c5, NULL, NULL) == FLASH_OK && fabs(c5 - cost) < 1e-12 * cost, "flash_points_set_file_async");
  }
  { /* frames */
    const double* fp[2] = {&cloud[0][0], &cloud[N / 2][0]};
    int64_t fc[2] = {N / 2, N - N / 2};
    double X2[2 * 14], c2[2], g2[2 * 14];
    memcpy(X2, x, sizeof x); memcpy(X2 + 14, x, sizeof x);
    CHECK(flash_frames_set(ctx, fp, fc, 2) == FLASH_OK && flash_num_frames(ctx) == 2 && flash_cost_grad_frames(ctx, X2, 1, 10.0, c2, g2) == FLASH_OK &&
          fabs(c2[0] + c2[1] - cost) < 1e-9 * cost, "flash_frames_set + flash_cost_grad_frames: the two halves add up");
    CHECK(flash_cost_grad(ctx, x, 1, 10.0, &cost, NULL, NULL) == FLASH_ERR_STATE, "flash_cost_grad on a batch of frames -> FLASH_ERR_STATE");
    CHECK(flash_frames_set_async(ctx, fp, fc, 2) == FLASH_OK && flash_points_swap(ctx) == FLASH_OK, "flash_frames_set_async");
  }
  { /* ray casting straight down onto the box: depth = camera height - top of the box */
    double cam[12] = {1, 0, 0, 0.1, 0, -1, 0, -0.2, 0, 0, -1, 2.0}, ray[3] = {0, 0, 1}, depth = 0, hit[3]; int64_t nh = 0;
    CHECK(flash_raycast(ctx, x, ray, 1, cam, &depth, hit, &nh) == FLASH_OK && nh == 1 && fabs(depth - (2.0 - 0.35)) < 1e-4 && fabs(hit[2] - 0.35) < 1e-4, "flash_raycast: depth %.6f", depth);
  }
  { /* on-device solver: recover a shifted box */
    flash_points_set(ctx, &cloud[0][0], N);
    double xs[14]; memcpy(xs, x, sizeof x); xs[4] += 0.03; xs[5] -= 0.02;
    flash_solver_params sp = {0.5, 0.05, 100, 0.0, 10.0};
    double cend = 0, c0 = 0; int32_t it = 0;
    flash_cost_grad(ctx, xs, 1, 10.0, &c0, NULL, NULL);
    CHECK(flash_estimate_state(ctx, xs, 1, &sp, NULL, &cend, &it) == FLASH_OK && it == 100 && cend < c0, "flash_estimate_state: cost %.6g -> %.6g in %d iterations", c0, cend, it);
  }
  { /* iso-surface of the box: count call, then fill */
    double lo[3] = {-0.4, -0.7, 0.1}, hi[3] = {0.6, 0.3, 0.5}; int64_t nt = 0;
    CHECK(flash_contour_mesh(ctx, x, 0, lo, hi, 0.02, 0.0, NULL, 0, &nt) == FLASH_OK && nt > 100, "flash_contour_mesh: %lld triangles", (long long)nt);
  }
  { /* device-resident points */
    CHECK(flash_points_set_device(ctx, (const void*)8, 4) == FLASH_ERR_INVALID, "flash_points_set_device: misaligned pointer rejected");
  }
  { /* communicator: a unique id can be made when NCCL is present; a world of one initialises */
    unsigned char id[FLASH_NCCL_UNIQUE_ID_BYTES];
    rc = flash_comm_unique_id(id);
    CHECK(rc == FLASH_OK || rc == FLASH_ERR_NCCL, "flash_comm_unique_id: %s", rc ? "NCCL not available (FLASH_ERR_NCCL)" : "ok");
    if (rc == FLASH_OK) CHECK(flash_comm_init(ctx, 0, 1, id) == FLASH_OK && flash_comm_world(ctx) == 1, "flash_comm_init(world = 1)");
  }
  flash_destroy(ctx);
  { /* grid mode */
    flash_options og = opt; og.sdf_mode = FLASH_SDF_GRID; og.grid_spacing = 0.01; og.grid_margin = 0.05;
    flash_ctx* cg = NULL;
    CHECK(flash_model_create(&d, &og, &cg) == FLASH_OK, "flash_model_create(grid mode)");
    if (cg) {
      int32_t dims[3]; double spacing = 0, err = 0;
      CHECK(flash_grid_report(cg, 0, dims, &spacing, &err) == FLASH_OK && dims[0] >= 61 && dims[0] <= 62 && fabs(spacing - 0.01) < 1e-12 && err < 0.01, "flash_grid_report: %d x %d x %d, max error %.2e m", dims[0], dims[1], dims[2], err);
      float* vals = (float*)malloc(sizeof(float) * dims[0] * dims[1] * dims[2]);
      double org[3], g2b[12];
      CHECK(flash_grid_download(cg, 0, org, g2b, vals) == FLASH_OK && fabs(vals[0] - sqrt(3 * 0.05 * 0.05)) < 1e-5, "flash_grid_download: corner value %.6f", vals[0]);
      free(vals);
      flash_destroy(cg);
    }
  }
  { /* deterministic mode and the single-process multi-device form (the same device twice is legal) */
    flash_options om = opt; om.deterministic = 1; om.n_devices = 2; om.devices[0] = 0; om.devices[1] = 0;
    flash_ctx* cm = NULL;
    CHECK(flash_model_create(&d, &om, &cm) == FLASH_OK && flash_num_devices(cm) == 2, "flash_model_create(n_devices = 2, deterministic)");
    if (cm) {
      double c1 = 0, c2 = 0, g1[14], g2[14];
      CHECK(flash_points_set(cm, &cloud[0][0], N) == FLASH_OK && flash_num_points(cm) == N && flash_cost_grad(cm, x, 1, 10.0, &c1, g1, NULL) == FLASH_OK &&
            flash_cost_grad(cm, x, 1, 10.0, &c2, g2, NULL) == FLASH_OK && c1 == c2 && memcmp(g1, g2, sizeof g1) == 0 && fabs(c1 - ref) < 1e-5 * ref, "multi-device cost: bit-reproducible, %.9g", c1);
      flash_destroy(cm);
    }
  }
  if (urdf) {
    const char* dirs[1] = {pkg};
    flash_ctx* cu = NULL;
    CHECK(flash_model_load_urdf(urdf, dirs, pkg ? 1 : 0, &opt, &cu) == FLASH_OK && flash_num_states(cu) == 6, "flash_model_load_urdf: IRB140, 6 states");
    flash_destroy(cu);
  }
  printf("%s\n", failures ? "FAILED" : "PASSED");
  return failures ? 1 : 0;
}
