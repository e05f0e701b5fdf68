/* b2ba -- B200-native bundle-adjustment engine: C ABI (the drop-in boundary).
 *
 * The reference (cteufel13/VisualOdometry) has NO optimiser API; the seam this
 * library plugs into is the tail of VisualOdometry._create_keyframe (src/modules/vo.py:252-288,
 * insertion point vo.py:286-288), and the data it exchanges is the state the reference holds:
 *   - camera pose  self.T_wc / keyframe["T_wc"] 4x4 f64 (vo.py:19,57,286); BA optimises
 *     T_cw = inv(T_wc) = [R|t] as used for projection (vo.py:260-261, frontend.py:128-139)
 *   - landmarks    self.map_points: id -> xyz f64 (vo.py:17,281)
 *   - observations keyframe["feats"]["keypoints"] (1,N,2) f32 + keyframe["ids"] (N,) (vo.py:53,57)
 *   - intrinsics   self.K 3x3 f64, zero distortion (vo.py:11,139; dataset_loader.py:52-54)
 * The Python binding a maintainer adds on the reference side is shown in INTEGRATION.md.
 *
 * Plain pointers and sizes only; no torch / C++ types.  All functions return 0 on success,
 * <0 on error (see B2BA_E_*); b2ba_solve returns >0 solver outcomes.  A handle owns one problem
 * (single or batched) with all iteration state resident on its device.  One handle <-> one CUDA
 * stream at a time; a handle is not thread-safe; distinct handles are independent.
 */
#ifndef B2BA_H_
#define B2BA_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2ba b2ba_t;

enum { B2BA_OK = 0, B2BA_E_INVALID = -1, B2BA_E_CUDA = -2, B2BA_E_NCCL = -3, B2BA_E_DUPLICATE = -4,
       B2BA_E_STATE = -5, B2BA_E_UNSUPPORTED = -6 };
/* solver outcomes (b2ba_summary.status; b2ba_solve returns the max over windows) */
enum { B2BA_MAX_ITERS = 1, B2BA_CONVERGED = 2, B2BA_NUMERICAL_FAILURE = 3 };
enum { B2BA_LOSS_TRIVIAL = 0, B2BA_LOSS_HUBER = 1 };
enum { B2BA_SOLVER_AUTO = 0, B2BA_SOLVER_WINDOW = 1, B2BA_SOLVER_PCG = 2,
       B2BA_SOLVER_DENSE = 3 /* large path, dense blocked Cholesky of S on the FP64 tensor cores */ };
enum { B2BA_MEM_HOST = 0, B2BA_MEM_DEVICE = 1 };

/* Spec a0 parameters. Zero tolerances disable the test (parity / bench mode). */
typedef struct b2ba_options {
  int32_t max_iters;      /* default 10 */
  int32_t loss;           /* B2BA_LOSS_* */
  double loss_scale;      /* Huber a, pixels (config.py:32,35 scale) */
  double z_min;           /* cheirality gate, config.py:31 min_depth = 1e-3 */
  double init_mu;         /* initial trust-region radius, 1e4 */
  double ftol, gtol, ptol;
  int32_t solver;         /* B2BA_SOLVER_* */
  int32_t pcg_max_iters;  /* 200 */
  double pcg_tol;         /* 1e-10 */
  int32_t flags;          /* bit0: record per-iteration history (b2ba_dump B2BA_D_HIST_*);
                             bit1: never split a window over a thread-block cluster (<= 18 windows per
                             handle are otherwise solved by one 8- or 16-CTA cluster each: the latency path);
                             bit2 / bit3 force kernel instantiations that the engine would not pick by itself
                             (256-thread CTAs on the batched small path / the portable 8-CTA cluster for <= 8
                             windows): same results up to summation order; used by the parity tests;
                             bit4: time the phases of the next solves with CUDA events (b2ba_get_timings) */
  int32_t pad_;
} b2ba_options;

typedef struct b2ba_summary {
  int32_t status, iterations, accepted, rejected, n_invalid, pcg_iters;
  double initial_cost, final_cost, final_mu;
} b2ba_summary;

/* Device times of the last set_problem / solve (SURVEY §5 "Metrics": per-phase ms).  Filled when options.flags bit 4
 * is set; phases are recorded with CUDA events on the solve's stream between the kernel groups of an LM iteration and
 * summed over the iterations (large path; the window path is ONE kernel, so only setup_ms / solve_ms apply there).
 * Host ranges of the same phases are emitted as NVTX ranges ("b2ba:<phase>") in every mode. */
enum { B2BA_T_LINEARISE = 0,   /* camera pass + landmark blocks + Z records (a1-a3) */
       B2BA_T_SCHUR = 1,       /* S blocks / owned block rows from the pair lists (a4) */
       B2BA_T_EXCHANGE = 2,    /* multi-GPU, banded / dense solvers: owned rows pushed into every rank's S + flags */
       B2BA_T_SOLVE = 3,       /* reduced camera system: PCG (incl. its in-kernel all-gathers) / banded / dense (a5-a6) */
       B2BA_T_BACKSUB = 4,     /* camera update, back-substitution, trial cost (a7) */
       B2BA_T_CONTROL = 5,     /* partial sums (multi-GPU: + their peer exchange), LM control, commit */
       B2BA_T_PHASES = 8 };
typedef struct b2ba_timings {
  double setup_ms;                 /* index build + structure (a8), device time of the last set_problem */
  double solve_ms;                 /* last b2ba_solve, first to last kernel */
  double phase_ms[B2BA_T_PHASES];  /* per phase, summed over the LM iterations */
  int32_t iterations_enqueued, pad_;
} b2ba_timings;

void b2ba_default_options(b2ba_options* o);

int b2ba_create(b2ba_t** out, int device, const b2ba_options* opts);
void b2ba_destroy(b2ba_t* h);
const char* b2ba_last_error(const b2ba_t* h);

/* One problem.  poses_cw: Nc x 12 row-major [R(9) | t(3)] (world->camera); points: Np x 3;
 * obs_cam / obs_pt: No int32; obs_uv: No x 2 float32 (u,v pixels); fixed_cam: Nc bytes (1 = gauge-
 * fixed); intr: fx, fy, cx, cy.  mem: B2BA_MEM_HOST or B2BA_MEM_DEVICE for ALL array arguments.
 * Inputs are copied into engine-owned SoA device arrays; caller pointers are not retained. */
int b2ba_set_problem(b2ba_t* h, int32_t Nc, int32_t Np, int64_t No, const double* poses_cw,
                     const double* points, const int32_t* obs_cam, const int32_t* obs_pt,
                     const float* obs_uv, const uint8_t* fixed_cam, const double intr[4], int mem);

/* B independent problems (C4), CSR-packed: cam_off/pt_off (B+1 int32), obs_off (B+1 int64) are
 * always HOST arrays; obs_cam/obs_pt are window-LOCAL indices. */
int b2ba_set_problem_batch(b2ba_t* h, int32_t B, const int32_t* cam_off, const int32_t* pt_off,
                           const int64_t* obs_off, const double* poses_cw, const double* points,
                           const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv,
                           const uint8_t* fixed_cam, const double intr[4], int mem);

/* Replace poses / points (same sizes) without rebuilding the index; NULL keeps the current. */
int b2ba_set_state(b2ba_t* h, const double* poses_cw, const double* points, int mem);
/* Restore the state given at set_problem / set_state (device-to-device). */
int b2ba_reset(b2ba_t* h);

/* Run LM for max_iters (<=0: options.max_iters) on `cuda_stream` (cudaStream_t, NULL = default).
 * summaries: HOST array of B entries (1 for a single problem) or NULL.  With NULL the call only
 * enqueues work (no synchronisation); otherwise it synchronises the stream and fills them. */
int b2ba_solve(b2ba_t* h, int32_t max_iters, b2ba_summary* summaries, void* cuda_stream);
int b2ba_get_summaries(b2ba_t* h, b2ba_summary* summaries);
/* Synchronises the last solve's stream; B2BA_E_STATE unless options.flags bit 4 was set for it. */
int b2ba_get_timings(b2ba_t* h, b2ba_timings* out);

int b2ba_get(b2ba_t* h, double* poses_cw, double* points, int mem);

/* Reprojection error of every observation at the CURRENT state (after b2ba_solve: the refined one), in pixels,
 * in the caller's observation order; -1 where the landmark is behind the cheirality gate.  The quantity the
 * reference gates new landmarks on (`frontend.py:139-142`: |projectPoints(X) - uv| < max_reproj_err, config.py:32),
 * here as the input of outlier culling after BA.  err: No doubles, HOST or DEVICE per `mem`. */
int b2ba_reproj_errors(b2ba_t* h, double* err, int mem);

/* Outlier culling after BA, on the device.  ONE kernel evaluates the reprojection error of every
 * observation at the current state, keeps it iff 0 <= err <= max_err -- the gate the reference applies to new landmarks
 * (`src/modules/frontend.py:139-142`, `src/config/config.py:31-32`; an observation that fails it corresponds to the
 * reference's "untracked" id -1, `src/modules/vo.py:53`, a landmark with no surviving observation to an entry
 * `_prune_map` would drop, `vo.py:35-47`) -- compacts the observation arrays in the caller's order (stable, single-pass
 * scan) and counts the survivors per landmark.  Outputs, any of which may be NULL: keep (No bytes, 1 = kept), live
 * (Np_tot int32: surviving observations per landmark), kept_index (the original index of every kept observation,
 * ascending; capacity No) -- HOST or DEVICE per `mem`; n_kept: ONE int64 on the HOST.
 * apply != 0: the handle's problem BECOMES the culled one -- compacted observation arrays, index rebuilt on the device,
 * state (and the state b2ba_reset restores) = the present one -- so optimise -> cull -> optimise needs no host round
 * trip of the observation arrays.  Landmarks keep their slots (one with live == 0 stays, unobserved). */
int b2ba_cull(b2ba_t* h, double max_err, uint8_t* keep, int32_t* live, int32_t* kept_index, int64_t* n_kept,
              int apply, int mem);

/* Batched two-view triangulation + gating: `triangulate_points` of the reference
 * (`src/modules/frontend.py:115-148`, called at `vo.py:262-270`).  P1, P2: K [R|t] of the two views (3 x 4 row-major,
 * HOST, the arguments of cv2.triangulatePoints at frontend.py:128-131); uv1, uv2: N x 2 float32 matched keypoints;
 * X: N x 3 (every match, like the reference's pts_3d before masking); keep[i] = depth in view 2 > min_depth &&
 * reprojection error in view 2 < max_reproj_err (frontend.py:135-143, config.py:31-32).  uv / X / keep HOST or DEVICE. */
int b2ba_triangulate(b2ba_t* h, int32_t N, const double P1[12], const double P2[12], const float* uv1,
                     const float* uv2, double min_depth, double max_reproj_err, double* X, uint8_t* keep, int mem);

/* Motion-only bundle adjustment: refine B camera poses, each against its own FIXED landmarks
 * -- the refinement the reference leaves to cv2.solvePnPRansac (src/modules/vo.py:131-150: X = map points of the
 * matched ids, uv = current keypoints, pose = inv(T_wc)).  obs_off (B+1, HOST, obs_off[0] = 0): problem b owns
 * observations [obs_off[b], obs_off[b+1]); points: No x 3 (the landmark of every observation); obs_uv: No x 2
 * float32; poses_cw: B x 12 [R | t], in/out; intr: fx, fy, cx, cy.  Same residual, robust loss, damping rule and
 * stopping tests as b2ba_solve (the handle's options), restricted to the 6 x 6 camera block.  Independent of the
 * problem set with b2ba_set_problem.  Returns the worst status (>0) or an error (<0); summaries (B, HOST) may be NULL. */
int b2ba_refine_poses(b2ba_t* h, int32_t B, const int64_t* obs_off, double* poses_cw, const double* points,
                      const float* obs_uv, const double intr[4], int32_t max_iters, b2ba_summary* summaries, int mem);

/* Robust pose tracking against FIXED landmarks from a pose prior: the GPU stand-in for the reference's
 * cv2.solvePnPRansac(X, uv, K, None, reprojectionError = pnp_reproj_err) call (src/modules/vo.py:131-150, config.py:35)
 * once the tracker has a previous pose.  `rounds` rounds of [motion-only LM on the current inlier set, Huber with
 * a = thr_r (the last of several rounds: trivial loss) -> every observation re-classified: inlier = depth >=
 * z_min and |reprojection error| < thr_r], thr_r = reproj_err * 2^(rounds - 2 - r) >= reproj_err (graduated; 6 rounds:
 * 16, 8, 4, 2, 1, 1 x reproj_err), which ends the way OpenCV's RANSAC does (consensus set by that threshold, pose
 * refined on it); no random sampling, so the result is deterministic.  pose_cw: 12 doubles [R | t], in (prior) /
 * out; points N x 3, obs_uv N x 2 float32; inliers: N bytes out; n_inliers (HOST) out.  pose / points / uv / inliers
 * HOST or DEVICE per `mem`.  Damping rule and stopping tests as b2ba_solve (the handle's options).  Returns the status of
 * the last round (>0) or an error (<0); fewer than 6 inliers after a round ends the call (the caller resets, as the
 * reference does when solvePnPRansac fails, vo.py:142-143). */
int b2ba_track_pose(b2ba_t* h, int64_t N, double* pose_cw, const double* points, const float* obs_uv,
                    const double intr[4], double reproj_err, int32_t rounds, int32_t iters_per_round,
                    uint8_t* inliers, int32_t* n_inliers, b2ba_summary* summary, int mem);

/* Sampled-consensus pose from 2D-3D matches: the GPU counterpart of the reference's
 * cv2.solvePnPRansac(X, uv, K, None, reprojectionError = pnp_reproj_err) call (src/modules/vo.py:135-141) for a tracker
 * that has a previous pose.  n_hyp hypotheses are evaluated at once: hypothesis j = the motion-only LM fit (hyp_iters
 * iterations, started from the prior pose_cw) of 4 distinct observations picked by a stateless hash of (seed, j);
 * every hypothesis is scored by its consensus (depth >= z_min and |reprojection error| < reproj_err over ALL N
 * observations); the largest consensus wins (ties: the lowest j) and is refined like OpenCV's RANSAC ends -- LM on the
 * consensus set, here two rounds (Huber a = reproj_err, then trivial loss; refine_iters iterations each) with the set
 * re-classified after each.  Deterministic for a given seed.  best_consensus (HOST, may be NULL): the winning
 * hypothesis's count before refinement; the other arguments as b2ba_track_pose.  N < 6: status 3, no inliers. */
int b2ba_ransac_pose(b2ba_t* h, int64_t N, double* pose_cw, const double* points, const float* obs_uv,
                     const double intr[4], double reproj_err, int32_t n_hyp, int32_t hyp_iters, uint64_t seed,
                     int32_t refine_iters, uint8_t* inliers, int32_t* n_inliers, int32_t* best_consensus,
                     b2ba_summary* summary, int mem);

/* Pipelined end-to-end batch solve (BASELINE config 4): set_problem_batch + solve + get in one call, with the
 * windows cut into `chunks` (<= 0: 8) contiguous chunks whose host->device copies, index build + solve and
 * device->host read-back overlap on three streams.  HOST buffers only (pinned memory makes the copies
 * asynchronous); window problems only (<= 24 cameras per window).  out_poses_cw / out_points / summaries (B
 * entries) receive the results; the handle keeps no problem afterwards (b2ba_get / b2ba_dump need set_problem).
 * Results are identical to set_problem_batch + solve + get (each window is solved by the same kernel). */
int b2ba_optimize_batch(b2ba_t* h, int32_t B, const int32_t* cam_off, const int32_t* pt_off,
                        const int64_t* obs_off, const double* poses_cw, const double* points,
                        const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv,
                        const uint8_t* fixed_cam, const double intr[4], int32_t max_iters,
                        double* out_poses_cw, double* out_points, b2ba_summary* summaries, int32_t chunks);

/* Per-stage intermediates for parity tests.  `what` selects the array; `bytes` is the size of dst
 * (HOST memory); returns the number of bytes written or <0. */
enum {
  B2BA_D_PM_PERM = 1, B2BA_D_PT_PTR, B2BA_D_CM_PERM, B2BA_D_CAM_PTR, B2BA_D_FIDX, B2BA_D_BLK_PTR,
  B2BA_D_BLK_COL,                       /* a8 index arrays (int32) */
  B2BA_D_SCHED,                         /* landmark schedule: per window, local ids by decreasing observation count (stable) */
  B2BA_D_RESIDUALS = 16, B2BA_D_WEIGHTS, B2BA_D_JC, B2BA_D_JP,   /* a1, original obs order */
  B2BA_D_B = 32, B2BA_D_GC, B2BA_D_C, B2BA_D_GP, B2BA_D_S, B2BA_D_GRED, B2BA_D_DC, B2BA_D_DP,
  B2BA_D_SCALARS,                       /* one LM iteration at init_mu from the CURRENT state */
  B2BA_D_HIST_COST = 64, B2BA_D_HIST_MU, B2BA_D_HIST_ACCEPT,     /* B x 64 */
  B2BA_D_POSE_HIST_COST = 80, B2BA_D_POSE_HIST_MU, B2BA_D_POSE_HIST_ACCEPT   /* b2ba_refine_poses: B x 64 */
};
int64_t b2ba_dump(b2ba_t* h, int what, void* dst, int64_t bytes);

/* Multi-GPU (C5 observation sharding): nccl_unique_id = the 128 bytes of an ncclUniqueId created by
 * rank 0 and broadcast by the caller (e.g. torch.distributed); must be called before set_problem.
 * Every rank passes the FULL problem to set_problem; the engine keeps its point range. */
int b2ba_create_comm(b2ba_t* h, const void* nccl_unique_id, int rank, int nranks);
/* Fills 128 bytes with a fresh ncclUniqueId (call on rank 0, broadcast, pass to b2ba_create_comm). */
int b2ba_nccl_unique_id(void* out128);
/* This rank's landmark range [p_lo, p_hi) of the NEXT set_problem (required when nranks > 1; ranges
 * of all ranks must tile [0, Np)).  The host side balances them by observation count. */
int b2ba_set_point_range(b2ba_t* h, int32_t p_lo, int32_t p_hi);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * The window solve (window_lm_kernel, the kernel b2ba_solve launches for problems of <= 24 cameras per window) with ALL
 * of its scratch -- the landmark ping-pong buffers and the per-landmark rows F | g_p | D_p -- in one file-scope
 * `static __device__` array of the library: the call allocates NOTHING (no cudaMalloc / cudaFree / stream-ordered
 * allocation, no events) and does not synchronise; it enqueues two memory operations and three kernels on `cuda_stream`.
 *   Set-up (may allocate, do it once, outside the no-allocation region): b2ba_create + b2ba_set_problem / _batch build the
 * index arrays the kernel walks; this call reads them and nothing else of the handle.
 *   poses_cw (Nc x 12 [R | t]) and points (Np x 3, the caller's landmark order) are DEVICE buffers of the caller, in: the
 * initial state, out: the refined one; summaries: DEVICE array of B entries (out).  max_iters <= 0: options.max_iters.
 *   Limit: the handle's problem may hold at most B2BA_NOALLOC_MAX_POINTS landmarks in total (144 bytes of scratch each: a
 * 4.5 MiB array that every GPU which loads the library holds until exit); more -> B2BA_E_UNSUPPORTED, use b2ba_solve.
 *   The array is shared by every call on a device and keeps the last call's contents (what must start at zero is cleared
 * by the call itself): calls that OVERLAP on the GPU -- two streams, two handles, two threads -- race; serialise them.
 *   The first call on a device resolves the array's device address (cudaGetSymbolAddress: no allocation, but make that
 * one call before a graph capture begins).  Results are those of b2ba_set_state + b2ba_solve + b2ba_get on the same
 * inputs (same kernel, same data); no history / timing options apply. */
#define B2BA_NOALLOC_MAX_POINTS 32768
int b2ba_solve_window_noalloc(b2ba_t* h, double* poses_cw, double* points, b2ba_summary* summaries, int32_t max_iters,
                              void* cuda_stream);

/* Measurement helper (bench.py roofline denominators; MEASURED_PEAKS.json has no fp64 entry): FP64 peak of `device`
 * in TFLOP/s with dependent-free DFMA chains and with mma.sync m8n8k4 f64 (DMMA), CUDA-event timed, ~50 ms. */
int b2ba_peak_flops(int device, double* dfma_tflops, double* dmma_tflops);

int b2ba_version(void);

#ifdef __cplusplus
}
#endif
#endif
