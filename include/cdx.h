/* cdx - C ABI of the B200-native GPIS + pregrasp hot path (drop-in for ComplianceDex's gpis.py /
 * optimize_pregrasp.py `prob` path).  The reference has NO FFI layer: its callers import plain Python classes
 * (optimize_pregrasp.py:5 `from gpis import GPIS`, :842-848 `optimizers[...]`).  This header is therefore the
 * boundary a maintainer binds with ctypes; each entry point cites the reference
 * function(s) it replaces.
 *
 * Conventions
 *  - every function returns 0 (CDX_OK) or a CDX_ERR_* code; cdx_last_error() gives the message (thread local).
 *  - pointers named *_dev are DEVICE pointers owned by the caller (float64, contiguous, row-major); pointers
 *    named *_host are host pointers and the call performs the H2D/D2H copies itself on `stream`.
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream). Calls are asynchronous on that
 *    stream unless stated; no internal threads; one cdx_ctx per device.
 *  - no CPU fallback exists: every compute entry point fails with CDX_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef CDX_H_
#define CDX_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cdx_ctx cdx_ctx;

enum { CDX_OK = 0, CDX_ERR_ARG = 1, CDX_ERR_CUDA = 2, CDX_ERR_NOT_SPD = 3, CDX_ERR_STATE = 4, CDX_ERR_ALLOC = 5 };
enum { CDX_KERNEL_TPS = 0, CDX_KERNEL_RBF = 1, CDX_KERNEL_JOINT = 2 }; /* gpis.py:9-14 */
enum { CDX_MAT_E11 = 0, CDX_MAT_CHOL = 1, CDX_MAT_CHOL_INV = 2, CDX_MAT_E11_INV = 3,
       CDX_VEC_ALPHA = 4, CDX_VEC_Y1 = 5, CDX_MAT_X1 = 6 };

const char* cdx_last_error(void);
int cdx_version(void);

/* one context per device; owns the factorised GPIS state and the pregrasp work buffers */
int cdx_create(int device, cdx_ctx** out);
int cdx_destroy(cdx_ctx* ctx);

/* ------------------------------------------------------------------ GPIS ------------------------------- */
/* GPIS.fit (gpis.py:33-40): R = max pairwise distance (tps/joint), y1 <- y1 - bias, E11 = k(X1,X1) + diag(noise^2);
 * then - what the reference defers to every pred/compute_normal call (LU at gpis.py:53, inverse at :77) - the
 * in-house blocked Cholesky E11 = L L^T, L^-1, E11^-1 and alpha = E11^-1 y1, once.
 * X1_dev [N,3], y1_dev [N] (labels BEFORE bias subtraction), noise_dev [N] or NULL (then noise_scalar is used). */
int cdx_gpis_fit(cdx_ctx* ctx, const double* X1_dev, const double* y1_dev, const double* noise_dev,
                 double noise_scalar, int64_t N, double bias, int kernel, double sigma, void* stream);

/* optional phase timing of cdx_gpis_fit / cdx_gpis_set_state (bench.py's fit workload): after cdx_gpis_fit_profile(ctx, 1)
 * a fit records CUDA events at its phase boundaries; cdx_gpis_fit_profile_read synchronises `stream` and returns the
 * milliseconds of {assembly (R, labels, E11), Cholesky, L^-1, E11^-1 = L^-T L^-1, alpha + packing} of the LAST fit. */
int cdx_gpis_fit_profile(cdx_ctx* ctx, int enable);
int cdx_gpis_fit_profile_read(cdx_ctx* ctx, double* ms_out5, void* stream);

/* GPIS.load_state_data (gpis.py:162-168): adopt a stored state (y1 already bias-subtracted, E11 as stored) and
 * factorise it. */
int cdx_gpis_set_state(cdx_ctx* ctx, const double* X1_dev, const double* y1_dev, const double* E11_dev, double R,
                       double bias, int64_t N, int kernel, double sigma, void* stream);

/* state read-back for GPIS.save_state_data (gpis.py:155-160) and for tests. `which` = CDX_MAT_* / CDX_VEC_*;
 * matrices are written dense [N,N] row-major (no padding). */
int cdx_gpis_get(cdx_ctx* ctx, int which, double* out_dev, void* stream);
/* synchronises `stream`; not_spd_col = -1 or the first failing pivot column of the Cholesky */
int cdx_gpis_info(cdx_ctx* ctx, int64_t* N, double* R, double* bias, double* prior_var, int* not_spd_col,
                  void* stream);

/* GPIS.pred (gpis.py:43-59) + the gradients the optimizers obtain by autograd through it
 * (optimize_pregrasp.py:733-736,758,763) + the un-normalised normal of GPIS.compute_normal (gpis.py:63-87):
 *   mean[M] = k(x,X1) alpha + bias,  grad_mean[M,3],  std[M] = sqrt|k(x,x) - k*^T E11^-1 k*|,  grad_std[M,3].
 * Any output may be NULL.  std alone uses the triangular factor (N^2 flop/query); grad_std uses E11^-1
 * (2 N^2 flop/query). */
int cdx_gpis_query(cdx_ctx* ctx, const double* X_dev, int64_t M, double* mean_dev, double* grad_mean_dev,
                   double* std_dev, double* grad_std_dev, void* stream);
/* which arithmetic evaluates std when grad_std is not requested (the contraction L^-1 k* is N^2 flop/query):
 *   -1  FP64 tensor MMA (mma.sync m8n8k4) always - bit-identical across batch sizes;
 *    0  auto: TPS states with 1024 <= padded N <= 8192 use the tcgen05 INT8 path (csrc/cdx_query_i8x.cu) with 5 byte slices
 *       (error-free integer products of a 38-bit fixed-point k* and L^-1; measured |d var| < 2e-8 prior), else FP64 MMA.  The choice depends on
 *       the state only, so a query's result does not depend on the batch it is part of;
 *  4..6 INT8 path with that many slices whenever the state is eligible (TPS, padded N <= 8192), any M. */
int cdx_gpis_set_var_mode(cdx_ctx* ctx, int mode);
/* kernel of the EXTENDED query (grad_std requested) for states of up to 448 padded training points: 0 (default) the
 * warp-specialised all-columns kernel (gpis_extw_kernel: k* generated once per query tile by generator warps, E11^-1 streamed
 * once by bulk TMA, DMMA warps synchronised by mbarriers), 2 its single-role predecessor (gpis_ext_kernel), 1 the
 * column-tiled kernel (gpis_var_kernel<ext>) that larger states use.  All meet the parity gates; w = E11^-1 k* is the same
 * bit for bit, the last bits of the outputs differ (summation order of the epilogue). */
int cdx_gpis_set_ext_mode(cdx_ctx* ctx, int mode);
/* launch shape of the tcgen05 INT8 path (same slices, same exact integer products, same summation tree per query):
 *   3 (default) a CTA PAIR issues tcgen05.mma.cta_group::2 (256 rows of L^-1 as the M dimension, 96 queries as N); each CTA
 *     generates the k* byte slices of half the queries into its own half of the B operand;
 *   1 a 2-CTA cluster of single-CTA MMAs that exchange the generated k* slices through distributed shared memory;
 *   2 no cluster (every CTA generates all of its block's k*). */
int cdx_gpis_set_i8_variant(cdx_ctx* ctx, int variant);
/* same through HOST buffers (copies on `stream`, synchronises before returning) - the e2e path of bench.py */
int cdx_gpis_query_host(cdx_ctx* ctx, const double* X_host, int64_t M, double* mean_host, double* grad_mean_host,
                        double* std_host, double* grad_std_host, void* stream);
/* mean / grad_mean with caller-supplied weights alpha_dev[N] (compute_normal's `index` subsets, gpis.py:69-80);
 * the bias is NOT added (gpis.py:80 differentiates mu without bias) */
int cdx_gpis_query_weights(cdx_ctx* ctx, const double* alpha_dev, const double* X_dev, int64_t M, double* mean_dev,
                           double* grad_mean_dev, void* stream);

/* GPIS.topcd (gpis.py:127-150): surface voxels of a [steps,steps,steps] float32 mean grid (row-major) - mean < 0 and at
 * least one 6-neighbour with mean >= 0, interior voxels only.  Writes the linear indices in ascending order to
 * idx_dev (capacity steps^3) and their number to count_dev[0]. */
int cdx_grid_surface(cdx_ctx* ctx, const float* mean_dev, int steps, long long* idx_dev, long long* count_dev, void* stream);
/* NO COUNTERPART IN THE REFERENCE - added for callers that cannot allocate (CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same arguments and results as cdx_grid_surface,
 * but the per-block counts live in a file-scope static __device__ array (256 KB per GPU that loads the library, held until
 * exit) instead of a stream-ordered allocation: the call allocates and frees nothing.  Limit: steps <= 256 (CDX_ERR_ARG
 * with a message beyond that).  All calls on one device share the array, so calls that overlap on the GPU race -
 * serialise them on one stream. */
int cdx_grid_surface_static(cdx_ctx* ctx, const float* mean_dev, int steps, long long* idx_dev, long long* count_dev,
                            void* stream);

/* ------------------------------------------------------------------ mesh distance (TorchSDF) ----------- */
/* compute_sdf (thirdparty/TorchSDF/torchsdf/sdf.py:34-64; binding csrc/bindings.cpp:22-27 ->
 * unbatched_triangle_distance_forward_cuda, kernel csrc/unbatched_triangle_distance_cuda.cu:176-254): points_dev [P,3],
 * face_vertices_dev [F,3,3], both of `dtype` (0 = float32, 1 = float64) -> dist_dev [P] SQUARED distance to the closest
 * triangle (rounded to float like the reference, :235), sign_dev [P] int32 (+1 on the side the face normal points to),
 * normals_dev [P,3] unit vector closest point -> point, closest_dev [P,3].  First face wins ties (:236,:243).  Any output
 * may be NULL.  F = 0 zero-fills the outputs (what the reference's pre-zeroed tensors keep). */
int cdx_sdf_forward(cdx_ctx* ctx, const void* points_dev, int64_t P, const void* face_vertices_dev, int64_t F, int dtype,
                    void* dist_dev, int32_t* sign_dev, void* normals_dev, void* closest_dev, void* stream);
/* search strategy of cdx_sdf_forward: 0 auto (bounding-sphere culled search for F >= 512 faces and P >= 128 x SMs points), 1 exhaustive point x face
 * scan (what the reference does), 2 culled always.  Both return the same face, bit for bit. */
int cdx_sdf_set_mode(cdx_ctx* ctx, int mode);
/* unbatched_triangle_distance_backward_cuda (:256-270): grad_points [P,3] = 2 grad_dist (points - closest) */
int cdx_sdf_backward(cdx_ctx* ctx, const void* grad_dist_dev, const void* points_dev, const void* closest_dev, int64_t P,
                     int dtype, void* grad_points_dev, void* stream);

/* ------------------------------------------------------------------ hand model + pregrasp -------------- */
/* Kinematic tree in URDF link order (robot_model.py:115-127, urdf_utils.py:28-83): body i has parent[i] < i
 * (-1 for the root), dof[i] = column of q driving it or -1 (fixed), axis[i] in {0,1,2} with axis_sign[i] = +-1
 * (rigid_body.py:149-154), fixed_rot[i] = (Rz(yaw) Ry(pitch)) Rx(roll) of the joint origin as the float32 3x3 the
 * reference builds (rigid_body.py:142), xyz[i] float32.  tips: 4 link ids + offsets (robot_model.py:255-263).
 * All HOST pointers.  The loaded hand's identity is kept on the context: loading the hand that is already there is free,
 * loading a different one first waits for all work of the device (the kernels of a running optimisation read it). */
int cdx_hand_load(cdx_ctx* ctx, int n_bodies, const int* parent, const int* dof, const int* axis,
                  const float* axis_sign, const float* fixed_rot, const float* xyz, int n_dofs, const int* tip_body,
                  const float* tip_offset);

/* ProbabilisticGraspOptimizer.forward_kinematics (optimize_pregrasp.py:657-669): q_dev [B,n_dofs] f64,
 * palm_dev [B,6] f64 (xyz + XYZ Euler) -> tips_dev [B,4,3] f64 (FK evaluated in float32 like the reference). */
int cdx_hand_fk(cdx_ctx* ctx, const double* q_dev, const double* palm_dev, int64_t B, double* tips_dev, void* stream);

/* KinGPISGraspOptimizer.forward_kinematics (optimize_pregrasp.py:431-438): compute_forward_kinematics(recursive=True) on
 * float64 angles (rigid_body.py:85-127: cos/sin in double, rounded into the float32 joint rotation; chain composed from the
 * tip towards the root) + palm_dev[b, 0:3].  palm_dev [B,6] f64 with zero Euler angles (the reference adds a constant
 * offset, no rotation).  -> tips_dev [B,4,3] f64. */
int cdx_hand_fk_recursive(cdx_ctx* ctx, const double* q_dev, const double* palm_dev, int64_t B, double* tips_dev,
                          void* stream);

/* ProbabilisticGraspOptimizer.compute_collision_loss (optimize_pregrasp.py:684-701) on world-frame anchor positions
 * anchor_dev [B,A,3] (cdx_hand_fk of the anchor links): sum over pairs of 1/dist where dist < threshold, + 0.1/z for
 * anchors below z = 0.02, + 1/z of the palm below 0.02 when palm_dev [B,6] is given (optimize_palm).  left_dev /
 * right_dev: int32 [npairs] DEVICE arrays of anchor indices.  cost_dev [B]. */
int cdx_collision_cost(cdx_ctx* ctx, const double* anchor_dev, int64_t B, int A, const int* left_dev, const int* right_dev,
                       int npairs, double threshold, const double* palm_dev, double* cost_dev, void* stream);

typedef struct cdx_pregrasp_params {
  int32_t num_copies;            /* len(pregrasp_coefficients), optimize_pregrasp.py:623-625 (<= 4) */
  float coeffs[4][4];            /* float32 as the reference stores them (:644) */
  float weights[4];              /* pregrasp_weights: float32 values, widened by .double() (:645) */
  float ref_q[16];               /* float32 (:634) */
  float box_lo[4][3], box_hi[4][3]; /* tip_bounding_box float32 (:643) */
  float com[3];                  /* COM as it lands in the float32 dummy_tip (:90-92) */
  float dummy_k;                 /* float32(gravity_const * mass / M), gravity_const = 10.0 (:94,:726) */
  float dummy_M;                 /* M = 2.0 (:73) */
  int32_t gravity;               /* 0: no dummy gravity finger (:98-99) */
  double uncertainty;            /* :630 */
  int32_t optimize_target, optimize_palm; /* :621,:631 */
  float cos_mu;                  /* torch.sqrt(1/(1+torch.tensor(friction_mu)**2)): float32 (:111,:707) */
  uint64_t noise_seed;           /* keyed stream replacing torch.rand_like at :61 (cdx_noise.cuh) */
  int64_t env_offset;            /* global index of candidate 0 (multi-GPU sharding) */
  double lr[5];                  /* q, compliance, target, palm pos, palm ori (:782-796) */
  double beta1, beta2, eps;      /* torch.optim.Adam defaults 0.9, 0.999, 1e-8 (:796) */
  double compliance_min;         /* 80.0 (:833) */
  int32_t track_after;           /* best-iterate tracking for s > 20 (:823); -1 = every iteration (Kin optimizers, :490) */
  int32_t variant;               /* 0: ProbabilisticGraspOptimizer (:614-839, Adam).  1: KinGPISGraspOptimizer (:408-511):
                                  * one copy with coefficient 1, reward x5, centre cost x10, force clamp 1, ref cost x20,
                                  * tar_dist x10, no contact-margin / pregrasp / palm terms, q used without the .float()
                                  * cast (:436), torch.optim.RMSprop (alpha below, eps), whole-batch margin latched at the
                                  * last improving iteration (:491-494).  2: GPISGraspOptimizer (:322-406): no hand - the
                                  * "q" slot holds tip_pose [B,12]; reward x25, variance cost = uncertainty * sum_f std,
                                  * no ref cost, tips clamped to the box (:400), target NOT clamped (:402), RMSprop */
  double rms_alpha;              /* torch.optim.RMSprop default 0.99 */
} cdx_pregrasp_params;

/* ProbabilisticGraspOptimizer.optimize setup (optimize_pregrasp.py:771-804): copies the initial iterate, zeroes
 * Adam moments and the best-iterate buffers. q0 [B,16], compliance0 [B,4], target0 [B,4,3], palm0 [B,6]. */
int cdx_pregrasp_init(cdx_ctx* ctx, const cdx_pregrasp_params* p, int64_t B, const double* q0_dev,
                      const double* compliance0_dev, const double* target0_dev, const double* palm0_dev, void* stream);
/* n_iters iterations of closure + best-tracking + Adam + clamps (optimize_pregrasp.py:805-834), no host sync.  The loop is
 * captured once per batch as a CUDA graph of up to 50 iterations and re-launched; iterations that do not
 * fill a chunk are launched directly. */
int cdx_pregrasp_run(cdx_ctx* ctx, int n_iters, void* stream);
/* launch tuning of cdx_pregrasp_run; results are bit-identical under every setting (tests/test_pregrasp_gpu.py).
 * groups: candidate groups pipelined on separate stream pairs, 0 = auto (2 from 768 candidates), 1..4.
 * occupancy: register allocation of the step kernel, 0 = auto by launch size, 2 = 255 registers, 3 = 168 registers.
 * use_graph: 0 = launch every iteration from the host. */
int cdx_pregrasp_set_tuning(cdx_ctx* ctx, int groups, int occupancy, int use_graph);
/* NaN policy of the reference (print and continue, optimize_pregrasp.py:819) without the per-iteration host sync: the step
 * kernel counts candidate-iterations whose total_loss was NaN since cdx_pregrasp_init.  Synchronises `stream`.
 * count_host[0] = that count, first_iter_host[0] = first iteration that saw one (-1: none; may be NULL). */
int cdx_pregrasp_nan_count(cdx_ctx* ctx, int64_t* count_host, int* first_iter_host, void* stream);
/* number of cdx_pregrasp_init calls this context has seen (0: none) and the current batch size: the pregrasp state is one
 * per context, so an optimizer object that shares a context checks that it still owns the state before reading it. */
int64_t cdx_pregrasp_generation(cdx_ctx* ctx, int64_t* B_out);
/* The read-back calls below take B = the number of candidates the caller's buffers were sized for; it must equal the
 * batch of the last cdx_pregrasp_init on this context (CDX_ERR_ARG otherwise: no overrun if another optimizer took the
 * context over with a different batch).
 * objective and its gradient at the CURRENT iterate without stepping (closure, :741-769) - for parity tests.
 * total_loss [B], margin [B,4], grads: q [B,16], compliance [B,4], target [B,4,3], palm [B,6]; any may be NULL */
int cdx_pregrasp_eval(cdx_ctx* ctx, int64_t B, double* total_loss_dev, double* margin_dev, double* gq_dev, double* gc_dev,
                      double* gt_dev, double* gp_dev, void* stream);
/* the reference's 5-tuple (:839) + opt_value (:802); any may be NULL */
int cdx_pregrasp_results(cdx_ctx* ctx, int64_t B, double* opt_q_dev, double* opt_compliance_dev, double* opt_target_dev,
                         double* opt_palm_dev, double* opt_margin_dev, double* opt_value_dev, void* stream);
/* current iterate (for tests) */
int cdx_pregrasp_state(cdx_ctx* ctx, int64_t B, double* q_dev, double* compliance_dev, double* target_dev, double* palm_dev,
                       void* stream);
/* KinGPISGraspOptimizer's `opt_margin` (:493): the WHOLE batch's margin [B,4] at the last iteration in which any candidate
 * improved, and that iteration's index (-1: never) in iter_dev[0] (int32).  flag = (opt_margin > 0).all() is the caller's. */
int cdx_pregrasp_latched_margin(cdx_ctx* ctx, int64_t B, double* margin_dev, int* iter_dev, void* stream);
/* local arg-min over opt_value with lowest-index tie-break: out_dev[0] = value, out_dev[1] = (double) GLOBAL index
 * (env_offset + local).  The cross-rank step is an all-gather of these 16 bytes. */
int cdx_pregrasp_argmin(cdx_ctx* ctx, double* out_dev, void* stream);

/* optional per-kernel timing (bench.py's roofline): after cdx_pregrasp_profile(ctx, 1) every cdx_pregrasp_run launches
 * directly (no graph) in the SAME candidate groups as the timed loop and records CUDA events around each group's launches;
 * cdx_pregrasp_profile_read synchronises `stream` and returns the summed milliseconds {wait for the overlapped mean query,
 * gpis extended query, gpis mean query, objective+Adam step} and the number of (iteration, group) launches covered, then
 * clears the record. */
int cdx_pregrasp_profile(cdx_ctx* ctx, int enable);
int cdx_pregrasp_profile_read(cdx_ctx* ctx, double* ms_out4, int* n_iters, void* stream);

/* how many cdx kernels this context has launched since creation (bench.py's gpu_launches) */
int64_t cdx_launch_count(cdx_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* CDX_H_ */
