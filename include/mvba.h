/*
 * mvba.h -- C ABI of the B200-native bundle-adjustment path of MultiView's rig_calibrator.
 *
 * The reference has no FFI layer: its solver boundary is the set of Ceres calls made by one
 * optimisation pass of rig_calibrator (reference rig_calibrator/bin/rig_calibrator.cc:1611-1919).
 * Every entry point below names the reference call site(s) it replaces.  A maintainer's binding
 * (the C++ glue that would sit in rig_calibrator.cc) is shown in INTEGRATION.md.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no exceptions cross this boundary.
 *  - Every call returns an int status: MVBA_OK (0) or a negative MVBA_E* code;
 *    mvba_last_error() gives a human readable message for the calling thread.
 *  - Parameter arrays are CALLER-OWNED host memory that must outlive the handle (exactly like the
 *    double* handed to ceres::Problem::AddResidualBlock).  mvba_evaluate and mvba_solve copy them
 *    host->device on entry; mvba_solve writes the optimised values back in place on exit.
 *  - A handle is not thread safe; distinct handles are independent.
 *  - All arithmetic is FP64.  Pixel and depth measurements are float32 at this boundary because
 *    the reference stores them as float (interest_point.cc:1839, dense_map_utils.cc:1382-1388).
 *  - There is NO CPU fallback: every compute entry point fails with MVBA_ENODEVICE when no
 *    sm_100 class CUDA device is usable.
 */
#ifndef MVBA_H_
#define MVBA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MVBA_ABI_VERSION 1

/* Block sizes, reference rig_calibrator/include/rig_calibrator/dense_map_utils.h (NUM_*_PARAMS). */
#define MVBA_NUM_RIGID_PARAMS 7   /* tx ty tz qx qy qz qw   (dense_map_utils.cc:159-178) */
#define MVBA_NUM_AFFINE_PARAMS 12 /* row-major top 3 rows   (dense_map_utils.cc:181-208) */
#define MVBA_NUM_XYZ_PARAMS 3
#define MVBA_MAX_DISTORTION 128   /* >= 2*58 = RPC degree 4 (rpc_distortion.cc:43-49)   */

enum {
  MVBA_OK = 0,
  MVBA_EINVAL = -1,    /* bad argument / inconsistent sizes                              */
  MVBA_ENODEVICE = -2, /* no usable CUDA device (there is no CPU fallback)               */
  MVBA_ECUDA = -3,     /* a CUDA runtime call failed                                     */
  MVBA_ESTATE = -4,    /* call order violated (e.g. solve before observations were set)  */
  MVBA_EBRACKET = -5,  /* interpolation fraction outside [0,1]: rig_calibrator.cc:346    */
  MVBA_ENCCL = -6,     /* NCCL failure / NCCL not loadable                               */
  MVBA_ENUMERIC = -7,  /* non finite cost at the initial point                           */
  MVBA_EPEER = -8      /* multi-GPU: another rank failed before a collective (its own call
                          returns the cause); every rank returns instead of blocking         */
};

/* Residual-block kinds, in the order the reference creates them for one track
 * (rig_calibrator.cc:1614-1903). */
enum {
  MVBA_BLOCK_CAM = 0,        /* BracketedCamError        rig_calibrator.cc:419-514  (2 rows) */
  MVBA_BLOCK_DEPTH_TRI = 1,  /* BracketedDepthError      rig_calibrator.cc:522-609  (3 rows) */
  MVBA_BLOCK_DEPTH_MESH = 2, /* BracketedDepthMeshError  rig_calibrator.cc:615-705  (3 rows) */
  MVBA_BLOCK_MESH_TRI = 3,   /* XYZError vs mesh         rig_calibrator.cc:1857-1883 (3 rows) */
  MVBA_BLOCK_TRI = 4         /* XYZError vs initial xyz  rig_calibrator.cc:1886-1901 (3 rows) */
};

typedef struct mvba_problem mvba_problem;

/* Static shape of one pass.  Replaces the flag-derived locals of main()
 * (rig_calibrator.cc:1401-1492). Sensor 0 is the reference sensor
 * (dense_map_utils.cc:940-1057 requires the ref sensor first). */
typedef struct {
  int32_t num_sensors;           /* T */
  int32_t num_ref_frames;        /* R  (world_to_ref_vec holds 7*R doubles)               */
  int32_t num_cameras;           /* C  (images after bracketing, cams.size())             */
  int32_t num_landmarks;         /* P  (tracks, pid_to_cid_fid.size()) -- local to a rank  */
  int32_t no_rig;                /* --no_rig                                              */
  int32_t affine_depth_to_image; /* --affine_depth_to_image (12-parameter depth block)    */
  double robust_threshold;       /* --robust_threshold: Cauchy scale of cam/depth/mesh     */
  double depth_tri_weight;       /* --depth_tri_weight   (rig_calibrator.cc:1755-1794)     */
  double depth_mesh_weight;      /* --depth_mesh_weight  (rig_calibrator.cc:1797-1840)     */
} mvba_config;

/* Pointers to the caller-owned parameter blocks; identical packing to the reference's
 * std::vector<double> storage (rig_calibrator.cc:1401-1492).  Pointer identity == block
 * identity, as with Ceres. Unused blocks may be NULL (world_to_cam unless no_rig;
 * depth_* when no depth residual exists). */
typedef struct {
  double* world_to_ref;     /* 7*R   world_to_ref_vec                                      */
  const double* ref_stamps; /* R     ref_timestamps (constants)                            */
  double* ref_to_cam;       /* 7*T   ref_to_cam_vec                                        */
  double* offsets;          /* T     ref_to_cam_timestamp_offsets                          */
  const double* offset_lo;  /* T     min_timestamp_offset (bounds, rig_calibrator.cc:1742) */
  const double* offset_hi;  /* T     max_timestamp_offset                                  */
  double* focal;            /* T     focal_lengths (one scalar per sensor, :1477)          */
  double* center;           /* 2*T   optical_centers                                       */
  double* distortion;       /* sum_t nd_t, sensor-major                                    */
  double* depth_to_image;   /* (7|12)*T depth_to_image_vec (unit-determinant part)         */
  double* depth_scale;      /* T     depth_to_image_scales                                 */
  double* world_to_cam;     /* 7*C   world_to_cam_vec (no_rig only)                        */
  double* xyz;              /* 3*P   xyz_vec                                               */
} mvba_params;

/* Which blocks float.  Mirrors the SetParameterBlockConstant calls at
 * rig_calibrator.cc:1702-1752, 1787-1793, 1834-1839.  Arrays have T entries (0 = constant). */
typedef struct {
  const uint8_t* focal;          /* "focal_length" in --intrinsics_to_float[t]             */
  const uint8_t* center;         /* "optical_center"                                       */
  const uint8_t* distortion;     /* "distortion" (ignored when nd_t == 0, :1709)           */
  const uint8_t* rig;            /* --rig_transforms_to_float (never for sensor 0)         */
  const uint8_t* depth_to_image; /* --depth_to_image_transforms_to_float                   */
  const uint8_t* poses;          /* --camera_poses_to_float; with a rig only poses[0]      */
                                 /* (the ref sensor) matters, :1719                        */
  int32_t float_timestamp_offsets; /* --float_timestamp_offsets                            */
  int32_t float_scale;             /* --float_scale (off when affine, :1787)               */
} mvba_float_mask;

/* ceres::Solver::Options subset the reference sets (rig_calibrator.cc:1910-1918) plus the Ceres
 * 1.14 defaults it inherits; mvba_solve_options_init fills exactly those values. */
typedef struct {
  int32_t max_num_iterations;    /* --num_iterations, 20                                   */
  double parameter_tolerance;    /* --parameter_tolerance, 1e-12                           */
  double function_tolerance;     /* 1e-16                                                  */
  double gradient_tolerance;     /* 1e-16                                                  */
  double initial_trust_region_radius; /* Ceres default 1e4                                  */
  double max_trust_region_radius;     /* 1e16                                               */
  double min_trust_region_radius;     /* 1e-32                                              */
  double min_relative_decrease;       /* 1e-3                                               */
  double min_lm_diagonal;             /* 1e-6                                               */
  double max_lm_diagonal;             /* 1e32                                               */
  double eta;                         /* inexact-Newton forcing term, 0.1                   */
  int32_t max_linear_solver_iterations; /* 500                                              */
  int32_t min_linear_solver_iterations; /* 0                                                */
  int32_t jacobi_scaling;             /* 1                                                  */
  int32_t linear_solver;              /* 0 = PCG on the reduced camera system (ITERATIVE_SCHUR),
                                         1 = dense FP64 tensor-core Cholesky of the reduced system */
  int32_t preconditioner;             /* 0 = block Jacobi of F'F+D, one block per parameter block (Ceres
                                         JACOBI, the reference's default); 1 = block Jacobi of S, same
                                         partition (SCHUR_JACOBI); 2 = block Jacobi of S with all shared
                                         per-sensor blocks (rig, intrinsics, depth) as one dense block */
  int32_t verbose;                    /* minimizer_progress_to_stdout                        */
} mvba_solve_options;

enum {
  MVBA_TERM_CONVERGENCE = 0,     /* parameter / function / gradient tolerance reached      */
  MVBA_TERM_NO_CONVERGENCE = 1,  /* max_num_iterations reached                              */
  MVBA_TERM_FAILURE = 2          /* trust region collapsed / numerical failure              */
};

typedef struct {
  double initial_cost;
  double final_cost;
  int32_t num_iterations;        /* LM iterations attempted (successful + unsuccessful)     */
  int32_t num_successful_steps;
  int32_t num_linear_iterations; /* total PCG iterations                                    */
  int32_t termination;
  int32_t num_jacobian_evals;    /* fused residual+Jacobian+normal-equation passes          */
  int32_t num_residual_evals;    /* cost-only passes                                        */
  int64_t num_kernel_launches;   /* CUDA kernels launched by this solve                     */
  double seconds_total;          /* host wall clock of mvba_solve incl. H2D/D2H             */
  double seconds_device;         /* CUDA-event time of the LM loop                          */
  double seconds_build;          /* CUDA-event time in the Jacobian/Schur build kernels     */
  double seconds_linear;         /* CUDA-event time in the linear solver                    */
  double seconds_update;         /* CUDA-event time in back-substitution + candidate cost   */
  double final_gradient_max_norm;
  double final_radius;
  double seconds_build_kernel;   /* CUDA-event time inside the fused build kernel launches only */
  int32_t num_build_launches;    /* launches of the fused build kernel (incl. the column-norm pass) */
  int32_t reserved;
} mvba_summary;

/* Per-iteration log, Ceres' progress table (minimizer_progress_to_stdout). */
typedef struct {
  double cost, cost_change, gradient_max_norm, step_norm, relative_decrease, radius;
  int32_t linear_iterations, step_accepted;
} mvba_iteration;

const char* mvba_last_error(void);
int mvba_abi_version(void);
/* Number of usable sm_100 devices; 0 when there is none (never an error). */
int mvba_device_count(void);

/* ceres::Problem problem;  (rig_calibrator.cc:1611).  device = CUDA ordinal. */
int mvba_create(const mvba_config* cfg, int device, mvba_problem** out);
/* ~ceres::Problem */
int mvba_destroy(mvba_problem* h);

/* Per-sensor constants that the cost functor copies from camera::CameraParameters
 * (rig_calibrator.cc:460, camera_params.cc:260-314): distorted image size (half size enters
 * Convert<UNDISTORTED_C,DISTORTED>) and the length of the distortion block
 * (0 none | 1 fisheye | 4,5 radtan | >5 RPC dist+undist halves). */
int mvba_set_sensor(mvba_problem* h, int32_t sensor, double image_width, double image_height,
                    int32_t num_distortion);

/* The double* arguments of every AddResidualBlock (rig_calibrator.cc:1695-1699, 1775-1781,
 * 1824-1829, 1874-1875, 1892-1893). */
int mvba_set_params(mvba_problem* h, const mvba_params* p);

/* cams[cid].{camera_type, beg_ref_index, end_ref_index, timestamp}
 * (camera_image.h:28-61; used at rig_calibrator.cc:1624-1666). */
int mvba_set_cameras(mvba_problem* h, int32_t num_cameras, const int32_t* sensor,
                     const int32_t* beg_ref, const int32_t* end_ref, const double* stamp);

/* The inlier (pid, cid, fid) triples in reference order: pid ascending, then cid ascending
 * (std::map iteration, rig_calibrator.cc:1614-1622).  pix = keypoint_vec[cid][fid].
 * has_depth/depth_xyz  = depthValue(cams[cid].depth_cloud, pix) result (:1757; may be NULL);
 * has_mesh/mesh_xyz    = pid_cid_fid_mesh_xyz[pid][cid][fid] != bad_xyz (:1801-1804; may be NULL).
 * Creates one CAM block per observation, plus DEPTH_TRI / DEPTH_MESH blocks as the reference
 * does, in the same order. */
int mvba_set_observations(mvba_problem* h, int64_t num_obs, const int32_t* pid, const int32_t* cid,
                          const float* pix, const uint8_t* has_depth, const float* depth_xyz,
                          const uint8_t* has_mesh, const double* mesh_xyz);

/* XYZError blocks (rig_calibrator.cc:1857-1901). kind = MVBA_BLOCK_MESH_TRI or MVBA_BLOCK_TRI.
 * pid ascending; a track without an inlier observation must not be listed (:1845-1855). */
int mvba_set_point_priors(mvba_problem* h, int32_t kind, int64_t n, const int32_t* pid,
                          const double* ref_xyz, double weight, double robust_threshold);

/* problem.SetParameterBlockConstant / SetParameter{Lower,Upper}Bound */
int mvba_set_float_mask(mvba_problem* h, const mvba_float_mask* m);

/* Number of scalar residuals / residual blocks in reference order. */
int mvba_num_residuals(mvba_problem* h, int64_t* num_residuals, int64_t* num_blocks);
/* For residual block b (reference order): kind[b], first residual index row0[b], and the index of
 * the originating observation or prior (src[b]).  pid_cid_fid_to_residual_index
 * (rig_calibrator.cc:1682) is row0 of the CAM blocks. Any output may be NULL. */
int mvba_residual_layout(mvba_problem* h, int32_t* kind, int64_t* row0, int64_t* src);

/* problem.Evaluate(eval_options, &cost, &residuals, NULL, NULL)  (rig_calibrator.cc:1163-1166).
 * apply_loss = eval_options.apply_loss_function.  residuals may be NULL. */
int mvba_evaluate(mvba_problem* h, int32_t apply_loss, double* cost, double* residuals);

/* Test hook: analytic Jacobian of `count` consecutive blocks of one kind (index within that
 * kind, creation order), all parameter blocks included whether floated or not, no loss
 * correction. Row-major [count][rows][width], columns in the reference's block order:
 *  CAM        (2 rows): beg7 end7 ref_to_cam7 xyz3 offset1 focal1 center2 distortion[nd_max]
 *  DEPTH_TRI  (3 rows): beg7 end7 ref_to_cam7 depth_to_image[7|12] scale1 xyz3 offset1
 *  DEPTH_MESH (3 rows): beg7 end7 ref_to_cam7 depth_to_image[7|12] scale1 offset1
 *  MESH_TRI/TRI (3 rows): xyz3
 * mvba_jacobian_width returns `width` for a kind. */
int mvba_jacobian_width(mvba_problem* h, int32_t kind, int32_t* width);
int mvba_evaluate_jacobian(mvba_problem* h, int32_t kind, int64_t first, int64_t count,
                           double* residuals, double* jacobian);

/* ceres::Solver::Options defaults as the reference sets them. */
int mvba_solve_options_init(mvba_solve_options* o);
/* ceres::Solve(options, &problem, &summary)  (rig_calibrator.cc:1919). Updates the parameter
 * arrays in place.  iterations (may be NULL) receives up to max_iterations_out records. */
int mvba_solve(mvba_problem* h, const mvba_solve_options* o, mvba_summary* s,
               mvba_iteration* iterations, int32_t max_iterations_out);

/* ---- per-track steps that run around every pass (device kernels, one thread per track) ----
 * The handle's observations are the features of the pass; `inlier` is one byte per observation in the
 * order they were given to mvba_set_observations (the reference's pid_cid_fid_inlier map,
 * rig_calibrator.cc:1528-1532).  A flag that is 0 on entry is skipped and stays 0. */
/* multiViewTriangulation (rig_calibrator/src/interest_point.cc:649-722): undistorts every inlier feature
 * (Convert<DISTORTED, UNDISTORTED_C>, camera_model/camera_params.cc:316-355), triangulates each track with
 * openMVG's 3-pass reweighted linear method (openMVG/multiview/triangulation_nview.cc:87-146) from the
 * CURRENT camera parameters and writes xyz[3 P]; tracks with fewer than two inlier rays or a non-finite
 * result lose all their features and keep the xyz they came in with. */
int mvba_triangulate(mvba_problem* h, uint8_t* inlier, double* xyz);
/* flagOutliersByTriAngleAndReprojErr (rig_calibrator/bin/rig_calibrator.cc:1045-1154) at the CURRENT
 * parameters: tracks whose largest ray convergence angle is below min_triangulation_angle (degrees) lose
 * all features, then features with a pixel residual norm above max_reprojection_error are dropped.
 * counts (may be NULL): features seen / removed by the angle step, seen / removed by the reprojection step. */
int mvba_flag_outliers(mvba_problem* h, double min_triangulation_angle, double max_reprojection_error,
                       uint8_t* inlier, int64_t counts[4]);
/* flagOutlierByExclusionDist (rig_calibrator.cc:1003-1038): features outside the centred crop box
 * (distorted_crop_size) of their sensor's image (image_size); both arrays hold 2 values per sensor. */
int mvba_flag_exclusion(mvba_problem* h, const double* image_size, const double* crop_size, uint8_t* inlier);

/* meshTriangulations / ray_mesh_intersect (rig_calibrator/src/texture_processing.cc:1436-1479, 1566-1625): the ray of
 * every inlier feature (undistorted pixel, CURRENT camera) against a triangle mesh (float32 vertices, 3 indices per
 * face; the reference's float32 triangle test of rayint/libs/acc/primitives.h), hits between min_ray_dist and
 * max_ray_dist.  has_mesh[N] / mesh_xyz[3 N] receive the per-feature result (no hit: 0 and bad_xyz = (1e100, 1e100,
 * 1e100), rig_calibrator.cc:1546) -- the arrays mvba_set_observations takes; pid_mesh_xyz[3 P] (may be NULL) the
 * per-track mean of the hits or bad_xyz, the reference points of the MESH_TRI priors. */
int mvba_mesh_intersect(mvba_problem* h, int64_t num_vertices, const float* vertices, int64_t num_faces,
                        const uint32_t* faces, double min_ray_dist, double max_ray_dist, const uint8_t* inlier,
                        uint8_t* has_mesh, double* mesh_xyz, double* pid_mesh_xyz);

/* fitRpcDist / fitRpcUndist (camera_model/rpc_distortion.cc:619-728; fitCurrDegRPC :556-617; what
 * CameraParameters::updateRpcUndistortion, camera_params.cc:214-254, and rig_calibrator/bin/fit_rpc.cc:80-146 run): fit an RPC
 * model (ratio of polynomials of degree rpc_degree, 1..6, in x and y) mapping in_xy[2 n] to out_xy[2 n] in the least-squares
 * sense, by degree continuation from the identity model, num_iterations Levenberg-Marquardt iterations per degree with
 * Ceres' rules and tolerances (function / gradient 1e-16, parameter_tolerance given).  jacobian_mode 0 follows the
 * reference's central-difference derivatives (closed form of the difference quotient), 1 uses exact derivatives.
 * coeffs receives 2 (rpc_degree + 1)(rpc_degree + 2) - 2 values in the reference's packed order (num_x, den_x, num_y, den_y);
 * stats (may be NULL): initial cost, final cost, LM iterations, successful steps.  Needs no problem handle. */
int mvba_rpc_fit(int32_t device, int64_t n, const double* in_xy, const double* out_xy, int32_t rpc_degree,
                 int32_t num_iterations, double parameter_tolerance, int32_t jacobian_mode, double* coeffs, double stats[4]);

/* mvba_rpc_fit for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  NO COUNTERPART IN THE REFERENCE.  Same arguments, same result
 * and same errors as mvba_rpc_fit, but it allocates nothing: every byte of device scratch (the staged samples, the per-CTA
 * partial normal equations and their folded sums) lives in one file-scope `static __device__` array of the library, about
 * 7.1 MB, which every GPU that loads libmvba.so holds until the process exits.  Limits: n <= MVBA_RPC_STATIC_MAX_N samples
 * (more: MVBA_EINVAL with a message naming the limit; use mvba_rpc_fit), and at most MVBA_RPC_STATIC_MAX_PARTS CTAs
 * accumulate (the allocating entry point uses up to two per SM; with n <= 64 MVBA_RPC_STATIC_MAX_PARTS both use the same
 * grid and return identical bits, beyond that the sums are folded in a different order and the fits agree to rounding).
 * No cudaMalloc, cudaFree, stream creation or allocating library call: the work runs on the calling thread's default
 * stream (cudaStreamPerThread) and the call returns after it has drained.  Every call shares the one array, which keeps
 * the previous call's contents (the entry point clears what must start at zero); calls that overlap on the GPU would race.
 * Calls from several threads of one process are serialised by a lock inside the library. */
#define MVBA_RPC_STATIC_MAX_N 65536
#define MVBA_RPC_STATIC_MAX_PARTS 96
int mvba_rpc_fit_static(int32_t device, int64_t n, const double* in_xy, const double* out_xy, int32_t rpc_degree,
                        int32_t num_iterations, double parameter_tolerance, int32_t jacobian_mode, double* coeffs, double stats[4]);

/* ---- measurement hooks (bench.py); device-resident inputs, CUDA-event timed ---- */
/* Uploads parameters and keeps them resident. */
int mvba_upload(mvba_problem* h);
/* One residual+Jacobian evaluation that STORES r and J for every CAM block (the unfused
 * north-star kernel (a)); returns CUDA-event milliseconds of `repeats` launches. */
int mvba_bench_jacobian(mvba_problem* h, int32_t repeats, float* ms_total);
/* One fused normal-equation build (Jacobian + J'J + Schur elimination); milliseconds. */
int mvba_bench_build(mvba_problem* h, int32_t repeats, float* ms_total);
/* Host wall clock of the one-time structure analysis behind the first evaluate / solve of a pass: the problem assembly
 * (what the AddResidualBlock loop of rig_calibrator.cc:1611-1903 costs the reference) and the reduced-system / tile tables. */
int mvba_host_seconds(mvba_problem* h, double* seconds_setup, double* seconds_prepare);
/* Which kernels the last mvba_solve planned for the normal-equation build (diagnostic; the tests use it to make sure a case
 * really runs the path it is meant to cover).  plan[0] pass A: 1 = per-image 6x6 accumulators, 0 = staged rows;
 * plan[1] pass B: 2 = staged window tiles (bulk-copied round tables), 1 = window tiles, 0 = none; plan[2] accumulator blocks
 * per warp of the tile kernel; plan[3] tracks per round; plan[4] tiles; plan[5] rounds; plan[6] tracks taking the generic
 * warp-per-track kernel; plan[7] padded local width of the widest window.  MVBA_ESTATE before the first solve. */
int mvba_build_plan(mvba_problem* h, int32_t plan[8]);

/* ---- multi-GPU: landmark sharding (each rank holds its own tracks; cameras replicated) ---- */
/* nccl_unique_id: 128 bytes from ncclGetUniqueId on rank 0 (see mvba_nccl_unique_id). */
int mvba_nccl_unique_id(uint8_t id128[128]);
int mvba_comm_init(mvba_problem* h, int32_t rank, int32_t nranks, const uint8_t id128[128]);

#ifdef __cplusplus
}
#endif
#endif /* MVBA_H_ */
