/*
 * flashsdf.h -- C ABI of libflashsdf, the B200 (sm_100a) implementation of the
 * tracking hot path of the Julia package `Flash`
 * (JuliaTagBot/point-cloud-signed-distance).
 *
 * The reference has no FFI of its own; its boundary is a set of Julia function
 * signatures.  Every entry point below names the Julia interface it replaces
 * (file:line under the reference tree).  The Julia-side `ccall` bindings are in
 * INTEGRATION.md and point-cloud-signed-distance_b200/julia/FlashCUDA.jl.
 *
 * Conventions
 *   - plain pointers and sizes only; all host buffers are caller-owned; the ctx
 *     owns device memory, its stream, events and (optionally) one NCCL communicator;
 *   - doubles at the ABI (the reference is Float64 end to end, src/models.jl:25,
 *     src/depthsensors.jl:16); fp32/fp64 mixing is internal to the kernels;
 *   - row-major everywhere: X is [C][nx], grad is [C][nx], points are [N][3]
 *     (the memory of a Julia Vector{SVector{3,Float64}});
 *   - every call returns 0 on success or a negative flash_status; the message is
 *     flash_last_error(ctx) (ctx may be NULL for creation failures).  Nothing
 *     throws across the ABI and nothing calls exit();
 *   - there is NO CPU fallback: without a CUDA device every compute entry point
 *     returns FLASH_ERR_NO_DEVICE.
 *   - a ctx is single-caller (the reference is single-threaded); distinct ctxs are
 *     independent.
 */
#ifndef FLASHSDF_H
#define FLASHSDF_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLASHSDF_ABI_VERSION 2

typedef struct flash_ctx flash_ctx;

typedef enum flash_status {
  FLASH_OK = 0,
  FLASH_ERR_INVALID = -1,    /* bad argument / malformed model description */
  FLASH_ERR_NO_DEVICE = -2,  /* no usable CUDA device (there is no CPU path) */
  FLASH_ERR_CUDA = -3,       /* a CUDA runtime call or kernel failed */
  FLASH_ERR_TOO_LARGE = -4,  /* model exceeds a compiled-in kernel limit */
  FLASH_ERR_STATE = -5,      /* call order (e.g. cost before points were set) */
  FLASH_ERR_NCCL = -6,       /* NCCL missing or a collective failed */
  FLASH_ERR_IO = -7          /* file ingestion failed */
} flash_status;

/* joint types: RigidBodyDynamics Fixed / Revolute(axis) / QuaternionFloating as used by
 * src/models.jl:26 (Revolute), :79 and :106 (QuaternionFloating), load_urdf :167. */
enum { FLASH_JOINT_FIXED = 0, FLASH_JOINT_REVOLUTE = 1, FLASH_JOINT_QUAT_FLOATING = 2 };

/* surface kinds: the BodyGeometry hierarchy, src/Flash.jl:35-48 */
enum { FLASH_SURF_CONVEX = 0, FLASH_SURF_RIGID_SKIN = 1, FLASH_SURF_DEFORMABLE_SKIN = 2 };

/* how convex bodies are evaluated */
enum { FLASH_SDF_EXACT = 0, FLASH_SDF_GRID = 1 };

/*
 * Flattened `Flash.Manipulator` (src/Flash.jl:62-65): a mechanism tree plus the ordered
 * `surfaces` vector.  Bodies are the moving bodies in tree order (a parent always precedes its
 * children; parent -1 is the world).  The configuration vector q concatenates the joints in body
 * order: fixed 0 values, revolute 1 (angle), quaternion-floating 7 ([w x y z tx ty tz], the
 * RigidBodyDynamics 0.0.2 layout).  The state vector is x = [q ; deformation_data]
 * (src/gradientdescent.jl:9-17), deformation_data holding 3 values per surface point of every
 * deformable skin, surfaces in order (src/Flash.jl:97-104).
 *
 * body transform:  T(body->world) = T(parent->world) * joint_to_parent * T_joint(q)
 * (body_to_joint is the identity in every builder of src/models.jl and for URDF input).
 *
 * Points of all surfaces are stored CSR style in `points`:
 *   convex surface s : points[point_offset[s] .. point_offset[s+1])  are the vertices, given in
 *                      the surface frame; surface_frame[s] maps that frame into body
 *                      surface_body[s] (the URDF visual origin, src/models.jl:154-164).  The body
 *                      is the convex hull of the vertices (EnhancedGJK semantics, src/models.jl:152).
 *   skin surface s   : the first n_surface_points[s] points are surface points (value 0), the
 *                      remaining ones skeleton points (value -1), the concatenation order of
 *                      src/Flash.jl:207-211; point_body[] gives the body each control point is
 *                      expressed in (a skin may span bodies: two_link_arm, src/models.jl:27-68).
 */
typedef struct flash_model_desc {
  int32_t abi_version;             /* FLASHSDF_ABI_VERSION */
  int32_t n_bodies;
  const int32_t* body_parent;      /* [n_bodies]      */
  const int32_t* joint_type;       /* [n_bodies]      */
  const double*  joint_axis;       /* [n_bodies][3]   revolute axis in the joint frame */
  const double*  joint_to_parent;  /* [n_bodies][12]  row-major [R|t], frameBefore -> parent body */
  int32_t n_surfaces;
  const int32_t* surface_kind;     /* [n_surfaces]    */
  const int32_t* surface_body;     /* [n_surfaces]    convex only (-1 = world) */
  const double*  surface_frame;    /* [n_surfaces][12] convex only */
  const int32_t* point_offset;     /* [n_surfaces+1]  */
  const int32_t* n_surface_points; /* [n_surfaces]    */
  const double*  points;           /* [n_points][3]   */
  const int32_t* point_body;       /* [n_points]      skins only (-1 = world) */
} flash_model_desc;

/*
 * Options.  The RBF switches exist because SpatialFields.jl (pinned 06046c2, REQUIRE.dev:23) is
 * not in the reference tree: `InterpolatingSurface(points, values, XCubed(), true)`
 * (src/Flash.jl:212) is restated as the cubic-RBF variational interpolant with a linear
 * polynomial, divided by its gradient norm -- the only definition consistent with
 * test/runtests.jl:17.  Defaults reproduce that.
 */
typedef struct flash_options {
  int32_t abi_version;
  int32_t device;                   /* CUDA device ordinal, -1 = current */
  int32_t rbf_normalize_by_gradient;/* 1 (default): skin value = f/|grad f| ; 0: raw f */
  int32_t rbf_polynomial_degree;    /* 1 (default): a0 + a.x ; 0: a0 only */
  int32_t sdf_mode;                 /* FLASH_SDF_EXACT (default) or FLASH_SDF_GRID */
  double  grid_spacing;             /* metres, grid mode; <=0 -> 0.004 */
  double  grid_margin;              /* metres around each hull's AABB covered by its grid; <=0 -> 0.10 */
  /* 1: run-to-run bit-identical results (the reference sums the points in one fixed order, src/gradientdescent.jl:32):
   * static group -> warp schedule, per-tile partial sums combined in tile order instead of fp64 atomics.  0 (default):
   * dynamic schedule + atomics, a few per cent faster, sums in a different order every run. */
  int32_t deterministic;
  /* Single-process multi-GPU context (the reference's estimate_state is called from ONE Julia process,
   * src/tracking.jl:8-27): n_devices > 1 makes flash_model_create build one sub-context per entry of devices[]
   * (CUDA ordinals), each with its own streams; the devices' sums are added by one kernel on devices[0] that reads
   * the others' accumulators over NVLink peer memory, in a fixed order (no NCCL in this mode).  flash_points_set /
   * flash_frames_set shard the cloud (frames) over the devices, flash_cost_grad / flash_estimate_state return the
   * global result, flash_raycast shards the rays.  0 or 1: one device (`device`). */
  int32_t n_devices;
  int32_t devices[16];
  int32_t reserved[6];
} flash_options;

void flash_options_default(flash_options* opt);

int flash_abi_version(void);
const char* flash_last_error(const flash_ctx* ctx);

/* `Manipulator(mechanism, surfaces)` from Flash.Models.* (src/models.jl:70,97,135,170,176):
 * Julia builds the model, flattens it once, and keeps the ctx for the model's lifetime. */
int flash_model_create(const flash_model_desc* desc, const flash_options* opt, flash_ctx** out);
void flash_destroy(flash_ctx* ctx);

/*
 * `load_urdf(filename; package_path)` (src/models.jl:154-171: parse_urdf + parse_urdf_visuals + extract_convex_surfaces)
 * without Julia: the kinematic tree breadth-first from the single root link (welded to the world), one convex surface
 * per <visual> mesh (binary or ASCII STL, OBJ) with the visual <origin> as its frame, vertices welded exactly,
 * `package://` names resolved against package_dirs (each either the parent of the package directory or the package
 * directory itself) and the URDF's own directory.  No CUDA device is needed to load; errors return FLASH_ERR_IO /
 * FLASH_ERR_INVALID with the message in flash_last_error(NULL).
 * The flash_urdf owns the arrays flash_urdf_desc points to; flash_model_create copies what it needs, so the flash_urdf
 * may be freed right after.  flash_urdf_set_joint_type replaces a body's joint (the notebooks swap the IRB140's base
 * for a quaternion-floating joint, examples/irb_and_squishable.ipynb:268-270).
 */
typedef struct flash_urdf flash_urdf;
int flash_urdf_load(const char* urdf_path, const char* const* package_dirs, int32_t n_package_dirs, flash_urdf** out);
const flash_model_desc* flash_urdf_desc(const flash_urdf* u);
const char* flash_urdf_body_name(const flash_urdf* u, int32_t body);        /* link names, in body order */
const char* flash_urdf_surface_mesh(const flash_urdf* u, int32_t surface);  /* resolved mesh file of a surface */
int flash_urdf_set_joint_type(flash_urdf* u, int32_t body, int32_t joint_type);
void flash_urdf_free(flash_urdf* u);
/* flash_urdf_load + flash_model_create in one call */
int flash_model_load_urdf(const char* urdf_path, const char* const* package_dirs, int32_t n_package_dirs,
                          const flash_options* opt, flash_ctx** out);

/* num_positions / num_states (src/Flash.jl:90) and the surface count */
int flash_num_positions(const flash_ctx* ctx);
int flash_num_states(const flash_ctx* ctx);
int flash_num_surfaces(const flash_ctx* ctx);

/* `CostFunctor(manipulator, sensed_points)` (src/gradientdescent.jl:41-47) captures the frame's
 * points once; so does this.  xyz is the memory of Vector{SVector{3,Float64}}.  Points are
 * rounded to fp32 on upload (stated in DESIGN.md). */
int flash_points_set(flash_ctx* ctx, const double* xyz, int64_t n);
int flash_points_set_f32(flash_ctx* ctx, const float* xyz, int64_t n);
/* zero-copy: a device buffer of n float4 {x,y,z,unused}, evaluated in the caller's order (no sort; group balls only).
 * The buffer must stay valid AND UNCHANGED until it is replaced by another flash_points_set* call: the group
 * balls computed here bound its contents, and the producer's work must be complete before the call. */
int flash_points_set_device(flash_ctx* ctx, const void* device_float4, int64_t n);
int64_t flash_num_points(const flash_ctx* ctx);

/*
 * Frame pipelining (the live loop of examples/irb_and_squishable.ipynb:495-503 evaluates frame k while frame k+1
 * arrives).  The ctx owns TWO point sets and a copy stream: flash_points_set_async uploads, Hilbert-sorts and groups
 * the next frame into the back set while kernels issued on the compute stream keep reading the front set;
 * flash_points_swap makes the back set current for every LATER call (stream-ordered, it does not block the host).
 * Pageable caller memory is staged into a pinned ring inside the ctx before the call returns; pinned caller memory
 * (cudaHostAlloc / cudaHostRegister) is read by DMA and must stay valid until the first compute call after the swap
 * has returned (or flash_points_sync).
 */
int flash_points_set_async(flash_ctx* ctx, const double* xyz, int64_t n);
int flash_points_set_async_f32(flash_ctx* ctx, const float* xyz, int64_t n);
int flash_points_swap(flash_ctx* ctx);
/*
 * `read_point_cloud(file)` (src/depthdata.jl:19-30): first line the camera origin `x, y, z`, then one return
 * `x, y, z, r, g, b` per line.  flash_read_point_cloud parses into caller buffers (xyz [cap][3], rgb [cap][3] or NULL;
 * *n receives the number of returns in the file, which may exceed cap: call with cap = 0 first to size the buffers).
 * flash_points_set_file / flash_points_set_file_async parse straight into the ctx's pinned staging ring and upload from
 * there (the async form fills the back set: flash_points_swap makes it current).  FLASH_ERR_IO on unreadable files.
 */
int flash_read_point_cloud(const char* path, double origin[3], double* xyz, double* rgb, int64_t cap, int64_t* n);
int flash_points_set_file(flash_ctx* ctx, const char* path, double origin[3]);
int flash_points_set_file_async(flash_ctx* ctx, const char* path, double origin[3]);
int flash_points_sync(flash_ctx* ctx);   /* host-blocks until every upload issued so far has finished */

/*
 * A batch of F frames (BASELINE.json configs[3]: 8 Kinect frames x 16 candidate configurations per frame).  Frame f is
 * xyz[f][counts[f]][3]; all frames are uploaded, sorted and grouped in one pass (F <= FLASH_MAX_FRAMES).
 * flash_cost_grad_frames then evaluates C candidate states PER FRAME in one launch: X is [F][C][nx], cost [F][C],
 * grad [F][C][nx] (may be NULL); state X[f][c] sees frame f only.
 */
#define FLASH_MAX_FRAMES 64
int flash_frames_set(flash_ctx* ctx, const double* const* xyz, const int64_t* counts, int32_t F);
int flash_frames_set_async(flash_ctx* ctx, const double* const* xyz, const int64_t* counts, int32_t F);
int flash_num_frames(const flash_ctx* ctx);
int flash_cost_grad_frames(flash_ctx* ctx, const double* X, int32_t C, double w_def, double* cost, double* grad);

/*
 * The hot path.  Replaces `(f::CostFunctor)(x)` + ForwardDiff's chunked gradient
 * (src/gradientdescent.jl:49-65, src/tracking.jl:16-21) for a batch of C candidate states:
 *   cost[c]  = sum_p skin_c(p)^2 + w_def * sum |deformation|^2     (RAW, not divided by N:
 *              src/tracking.jl:18-20 divides and calls the callback on the Julia side)
 *   grad[c]  = d cost[c] / d X[c]   including the in-cost quaternion normalisation
 *              (src/gradientdescent.jl:30) -- may be NULL
 *   argmin   = [C][N] index into `surfaces` of the first minimum (src/Flash.jl:267) -- may be NULL
 * With a communicator attached (flash_comm_init) the points of all ranks are summed with one
 * ncclAllReduce before the gradient epilogue, so every rank returns the global cost/gradient.
 */
int flash_cost_grad(flash_ctx* ctx, const double* X, int32_t C, double w_def,
                    double* cost, double* grad, int32_t* argmin);

/* `Flash.skin(state)(x)` (src/Flash.jl:265-268) for n query points under one state x:
 * dist[n] (and optionally the argmin surface index). */
int flash_skin_eval(flash_ctx* ctx, const double* x, const double* xyz, int64_t n,
                    double* dist, int32_t* argmin);

/*
 * flash_skin_eval for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * flash_skin_eval, but the call allocates nothing, on the device or on the host side of the CUDA
 * runtime: every device temporary (the state, the kinematics / RBF tables prepare_kernel
 * derives from it, the query points as doubles and as float4, the distances, the argmin, the
 * status word) lives in ONE file-scope `static __device__` array of
 * FLASH_STATIC_SCRATCH_BYTES (4 MiB) that every GPU which loads the library holds until exit.
 * Limits, checked before anything is launched (FLASH_ERR_TOO_LARGE with a message naming them):
 *   n <= FLASH_STATIC_MAX_QUERY (65,536 query points per call), and
 *   8 * (num_states + the model's table lengths) <= FLASH_STATIC_STATE_BYTES (512 KiB).
 * x, xyz, dist and argmin may be host or device pointers (copied with cudaMemcpyDefault).
 * The array is shared by every context on a device and keeps the last call's contents; the call
 * clears what must start at zero (the status word), holds a process-wide lock and returns only
 * when its stream has drained, so calls made through this entry point never overlap -- any two
 * users of the array that DID overlap on the GPU would race.  A multi-device context
 * evaluates on its first device, as flash_skin_eval does.
 */
#define FLASH_STATIC_MAX_QUERY 65536
#define FLASH_STATIC_STATE_BYTES (512 * 1024)
#define FLASH_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int flash_skin_eval_static(flash_ctx* ctx, const double* x, const double* xyz, int64_t n,
                           double* dist, int32_t* argmin);

/*
 * `DepthSensors.raycast_depths` / `raycast_points` (src/depthsensors.jl:56-113).
 * rays [R][3] are the SENSOR-frame unit rays (generateKinectRays, :10-30) in row-major order,
 * cam[12] the row-major [R|t] sensor->world transform.  depths[R] receives the marching distance
 * or NaN for a miss; hits (may be NULL) receives the world points of the non-NaN rays in ray
 * order, *nhits their number.
 */
int flash_raycast(flash_ctx* ctx, const double* x, const double* rays, int64_t R,
                  const double cam[12], double* depths, double* hits, int64_t* nhits);

/* generateKinectRays(rows, cols, vertical_fov, horizontal_fov), src/depthsensors.jl:10-30;
 * host-side helper writing rays[rows*cols][3] row-major (v outer, u inner). */
int flash_kinect_rays(int32_t rows, int32_t cols, double vertical_fov, double horizontal_fov,
                      double* rays);

/*
 * `estimate_state` with a NaiveSolver (src/tracking.jl:8-27) run as an on-device loop for a batch
 * of C warm starts: per iteration g = grad(cost/N); stop when |g|_2 < grad_tol;
 * step = -rate * g ./ precondition_divisors, rescaled to |step|_2 <= max_step; x += step.
 * (The step rule is SimpleGradientDescent's, which is not in the reference tree -- inferred from
 * the keyword names at src/tracking.jl:12-15 and examples/irb140.ipynb cell 9.)
 * divisors may be NULL (all ones).  X is updated in place; cost_out[C] (may be NULL) receives
 * the raw cost at the returned x; iters_out[C] (may be NULL) the iterations taken.
 */
typedef struct flash_solver_params {
  double rate;
  double max_step;
  int32_t iteration_limit;
  double gradient_convergence_tolerance; /* <=0: never stop early */
  double deformation_cost_weight;        /* src/gradientdescent.jl:7, default 10 */
} flash_solver_params;

int flash_estimate_state(flash_ctx* ctx, double* X, int32_t C, const flash_solver_params* p,
                         const double* precondition_divisors, double* cost_out, int32_t* iters_out);

/* ---- multi-GPU: one process per GPU, points sharded, one allreduce per call ---- */
#define FLASH_NCCL_UNIQUE_ID_BYTES 128
int flash_comm_unique_id(void* id128);             /* rank 0 creates, caller broadcasts */
int flash_comm_init(flash_ctx* ctx, int32_t rank, int32_t world, const void* id128);
int flash_comm_world(const flash_ctx* ctx);
/* devices of a single-process multi-GPU context (flash_options.n_devices); 1 for an ordinary context */
int flash_num_devices(const flash_ctx* ctx);

/* ---- measurement hooks (CUDA events on the ctx stream) ---- */
typedef struct flash_timing {
  float ms_prepare;    /* FK + RBF solve kernel            */
  float ms_main;       /* cost/gradient accumulation kernel */
  float ms_allreduce;  /* NCCL (0 without a communicator)   */
  float ms_epilogue;   /* adjoint solve + joint projection  */
  float ms_total;      /* first launch -> last kernel end   */
  int32_t launches;    /* kernels launched by the last call */
  float ms_group;      /* group pre-pass kernel (part of ms_main) */
} flash_timing;
int flash_profiling(flash_ctx* ctx, int32_t enable);
int flash_last_timing(const flash_ctx* ctx, flash_timing* t);
/* total kernels launched by this ctx since creation (bench.py's gpu_launches) */
int64_t flash_launch_count(const flash_ctx* ctx);
/* device properties the bench reports: SM count, clock (kHz), L2 bytes */
int flash_device_info(const flash_ctx* ctx, int32_t* sm_count, int32_t* clock_khz, int64_t* l2_bytes);
/* grid mode: max |grid - exact| sampled at bake time per surface (metres); -1 if not grid */
int flash_grid_report(const flash_ctx* ctx, int32_t surface, int32_t dims[3], double* spacing,
                      double* max_abs_err);

/* grid mode: copies surface's grid (x fastest, then y, then z; dims from flash_grid_report) and its
 * origin in the geometry frame (the surface frame re-centred on the hull centroid, returned as the
 * row-major [R|t] geometry->body transform) so that a caller can re-evaluate the same trilinear
 * formula -- the parity definition of grid mode. */
int flash_grid_download(const flash_ctx* ctx, int32_t surface, double origin[3], double geometry_to_body[12],
                        float* values);

/*
 * Headless replacement of the skin drawing in `draw` (src/Flash.jl:297-329: per surface
 * `contour_mesh(surface, lb, ub, 0.0, ...)` sent to a remote viewer): samples the field of `surface`
 * (an index into the model's surfaces, or -1 for skin(state) = the min over all of them) on the
 * regular grid lo + k*spacing (k = 0 .. ceil((hi-lo)/spacing)) in the world frame with the same
 * evaluators as the cost path, and extracts the iso-surface {field = iso} by marching tetrahedra.
 * triangles [cap][9] (may be NULL) receives up to cap triangles as three xyz vertices, oriented with the
 * normal towards increasing field (outwards); *n_triangles the number found, which can exceed cap
 * (call again with a larger buffer).  The triangle order is unspecified.
 */
int flash_contour_mesh(flash_ctx* ctx, const double* x, int32_t surface, const double lo[3], const double hi[3],
                       double spacing, double iso, float* triangles, int64_t cap, int64_t* n_triangles);

#ifdef __cplusplus
}
#endif
#endif /* FLASHSDF_H */
