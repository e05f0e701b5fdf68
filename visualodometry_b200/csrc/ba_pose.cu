// Motion-only bundle adjustment: refine ONE camera pose against fixed landmarks, the step
// the reference leaves to cv2.solvePnPRansac's internal refinement (`src/modules/vo.py:131-150`).  Same
// projection / residual / robust-loss / Jacobian code and the same Levenberg-Marquardt rule as the full
// problem (Spec a0) restricted to a single 6 x 6 camera block: H = sum w J_c^T J_c, g = sum w J_c^T r,
// D = clamp(diag H), (H + D/mu) delta = -g, left retraction, step quality against the model decrease.
//
// One CTA per pose problem, the whole LM loop in one launch: observations are strided over the CTA, the 21 + 6
// + 2 sums are reduced by a fixed shuffle tree + fixed-order cross-warp sum (no atomics, deterministic), thread
// 0 factors the damped 6 x 6 system, every thread evaluates the trial cost of its observations.
#include "ba_pose.cuh"

namespace b2 {
namespace {

constexpr int kPoseThreads = 128;
constexpr int kPoseWarps = kPoseThreads / 32;
constexpr int kNV = 29;   // B_c upper (21) | g_c (6) | cost | invalid count

__device__ __forceinline__ int packed21(int r, int s) { return r * 6 - r * (r - 1) / 2 + (s - r); }

// sum of v[0..n) over the CTA into out[0..n) (shared), identical on every thread afterwards
template <int N>
__device__ __forceinline__ void cta_sum(double* v, double* red, double* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    const double s = warp_sum(v[k]);
    if (lane == 0) red[warp * N + k] = s;
  }
  __syncthreads();
  if (threadIdx.x < N) {
    double s = 0.0;
    const int nw = (int)blockDim.x >> 5;               // 4 warps per pose problem, 1 for the minimal-sample fits
    for (int w = 0; w < nw; ++w) s += red[w * N + threadIdx.x];
    out[threadIdx.x] = s;
  }
  __syncthreads();
}

// the LM loop of problem b by one CTA; mask (per observation, or null) switches observations off
__device__ void pose_lm_run(const PoseData& d, const Params& P, int b, const unsigned char* __restrict__ mask) {
  __shared__ double red[kPoseWarps * kNV], lin[kNV], tri[2], pose[12], ptri[12], dl[8];
  __shared__ int ctl[2];
  const int tid = threadIdx.x;
  const long long o0 = d.obs_off[b], o1 = d.obs_off[b + 1];
  if (tid < 12) pose[tid] = d.poses[12 * (size_t)b + tid];
  __syncthreads();

  double mu = P.init_mu, nu = 2.0, cost = 0.0, initial_cost = 0.0;
  int n_acc = 0, n_rej = 0, ninv = 0, status = B2BA_MAX_ITERS, it = 0;
  bool have_lin = false;
  for (; it < P.max_iters; ++it) {
    const bool fresh = !have_lin;
    if (!have_lin) {
      // ---- linearise: B_c, g_c, cost, invalid count
      double v[kNV];
#pragma unroll
      for (int k = 0; k < kNV; ++k) v[k] = 0.0;
      for (long long i = o0 + tid; i < o1; i += blockDim.x) {
        if (mask != nullptr && !mask[i]) continue;
        const double* X = d.points + 3 * (size_t)i;
        const float2 uv = d.obs_uv[i];
        Proj o;
        project(P, pose, X[0], X[1], X[2], uv.x, uv.y, o);
        double jc[12];
        jac_cam(P, o, jc);
#pragma unroll
        for (int r = 0; r < 6; ++r) {
#pragma unroll
          for (int s = r; s < 6; ++s) v[packed21(r, s)] += o.w * (jc[r] * jc[s] + jc[6 + r] * jc[6 + s]);
          v[21 + r] += o.w * (jc[r] * o.r0 + jc[6 + r] * o.r1);
        }
        v[27] += 0.5 * o.rho;
        v[28] += o.valid ? 0.0 : 1.0;
      }
      cta_sum<kNV>(v, red, lin);
      cost = lin[27];
      ninv = (int)lin[28];
      if (it == 0) initial_cost = cost;
      have_lin = true;
    }
    // ---- thread 0: damped 6 x 6 Cholesky, delta, model-decrease terms
    if (tid == 0) {
      const double imu = 1.0 / mu;
      double A[6][6], D[6], g[6], x[6];
      bool ok = true;
      double gm = 0.0;
      for (int r = 0; r < 6; ++r) {
        D[r] = clampd(lin[packed21(r, r)]);
        g[r] = lin[21 + r];
        gm = fmax(gm, fabs(g[r]));
        for (int s = r; s < 6; ++s) A[s][r] = lin[packed21(r, s)] + (r == s ? D[r] * imu : 0.0);
      }
      for (int c = 0; c < 6 && ok; ++c) {              // A = L L^T, lower, in place
        double dj = A[c][c];
        for (int k = 0; k < c; ++k) dj -= A[c][k] * A[c][k];
        if (!(dj > 0.0) || !isfinite(dj)) { ok = false; break; }
        const double l = sqrt(dj);
        A[c][c] = l;
        for (int r = c + 1; r < 6; ++r) {
          double t = A[r][c];
          for (int k = 0; k < c; ++k) t -= A[r][k] * A[c][k];
          A[r][c] = t / l;
        }
      }
      double dDd = 0.0, gd = 0.0;
      if (ok) {
        for (int r = 0; r < 6; ++r) { double t = -g[r]; for (int k = 0; k < r; ++k) t -= A[r][k] * x[k]; x[r] = t / A[r][r]; }
        for (int r = 5; r >= 0; --r) { double t = x[r]; for (int k = r + 1; k < 6; ++k) t -= A[k][r] * x[k]; x[r] = t / A[r][r]; }
        for (int r = 0; r < 6; ++r) {
          ok = ok && isfinite(x[r]);
          dDd += x[r] * x[r] * D[r];
          gd += x[r] * g[r];
          dl[r] = x[r];
        }
      }
      dl[6] = 0.5 * (dDd * imu - gd);                  // model decrease
      dl[7] = gm;
      ctl[0] = ok ? 1 : 0;
      if (ok) retract_pose(pose, x, ptri);
    }
    __syncthreads();
    const bool step_ok = ctl[0] != 0;
    // ---- trial cost with the retracted pose
    double tv[2] = {0.0, 0.0};
    if (step_ok) {
      for (long long i = o0 + tid; i < o1; i += blockDim.x) {
        if (mask != nullptr && !mask[i]) continue;
        const double* X = d.points + 3 * (size_t)i;
        const float2 uv = d.obs_uv[i];
        Proj o;
        project(P, ptri, X[0], X[1], X[2], uv.x, uv.y, o);
        tv[0] += 0.5 * o.rho;
        tv[1] += o.valid ? 0.0 : 1.0;
      }
    }
    cta_sum<2>(tv, red, tri);
    // ---- LM control (identical on every thread: all inputs are in shared memory)
    const double tc = tri[0], model = dl[6], gm = dl[7];
    int accept = 0, stop = 0;
    double rho = 0.0;
    if (step_ok && model > 0.0) {
      rho = (cost - tc) / model;
      // a step that pushes MORE observations behind the cheirality gate (they cost 0) is rejected like a failed one
      accept = (isfinite(tc) && rho > kMinRatio && (int)tri[1] <= ninv) ? 1 : 0;
    }
    if (P.gtol > 0.0 && fresh && gm <= P.gtol) { status = B2BA_CONVERGED; break; }   // before taking the step
    if (accept) {
      if (P.ftol > 0.0 && fabs(cost - tc) <= P.ftol * cost) stop = B2BA_CONVERGED;
      double dn2 = 0.0, xn2 = 0.0;
      for (int r = 0; r < 6; ++r) dn2 += dl[r] * dl[r];
      for (int r = 0; r < 12; ++r) xn2 += pose[r] * pose[r];
      if (P.ptol > 0.0 && sqrt(dn2) <= P.ptol * (sqrt(xn2) + P.ptol)) stop = B2BA_CONVERGED;
    }
    const double mu_used = mu;
    __syncthreads();                                   // everyone has read pose / dl before they change
    if (accept) {
      const double t = 2.0 * rho - 1.0;
      mu = fmin(kMuMax, mu / fmax(1.0 / 3.0, 1.0 - t * t * t));
      nu = 2.0;
      cost = tc;
      ninv = (int)tri[1];
      if (tid < 12) pose[tid] = ptri[tid];
      have_lin = false;
      ++n_acc;
    } else {
      mu = mu / nu;
      nu *= 2.0;
      ++n_rej;
    }
    if (tid == 0 && d.hist_cost != nullptr && it < kPoseHist) {
      d.hist_cost[(size_t)b * kPoseHist + it] = cost;
      d.hist_mu[(size_t)b * kPoseHist + it] = mu_used;
      d.hist_accept[(size_t)b * kPoseHist + it] = accept;
    }
    if (!stop && !(mu > kMuMin)) stop = B2BA_NUMERICAL_FAILURE;
    __syncthreads();
    if (stop) { status = stop; ++it; break; }
  }
  if (tid < 12) d.poses[12 * (size_t)b + tid] = pose[tid];
  if (tid == 0) {
    b2ba_summary s;
    s.status = status; s.iterations = it; s.accepted = n_acc; s.rejected = n_rej; s.n_invalid = ninv; s.pcg_iters = 0;
    s.initial_cost = initial_cost; s.final_cost = cost; s.final_mu = mu;
    d.summaries[b] = s;
  }
  __syncthreads();                                     // the shared state may be reused by another run
}

__global__ void __launch_bounds__(kPoseThreads) pose_lm_kernel(PoseData d, Params P) {
  pose_lm_run(d, P, blockIdx.x, nullptr);
}

// the same loop by ONE warp per problem: the minimal-sample fits of the sampled consensus (4 observations each) have no
// use for more lanes, and 1024 of them fit the GPU in one wave
__global__ void __launch_bounds__(32) pose_lm_warp_kernel(PoseData d, Params P) {
  pose_lm_run(d, P, blockIdx.x, nullptr);
}

// Robust pose tracking: the GPU stand-in for the cv2.solvePnPRansac call of the reference's tracker
// (`src/modules/vo.py:131-150`) when a pose prior exists (the previous frame's pose).  `rounds` rounds of
//   motion-only LM on the current inlier set (Huber with a = thr_r; the last of several rounds: trivial loss)
//   -> every observation re-classified: inlier = in front of the camera and |reprojection error| < thr_r
// with thr_r = thr x 2^(rounds - 2 - r) >= thr (graduated: the last two rounds run at thr itself)
// (thr = the reference's `pnp_reproj_err`, `config.py:35`; OpenCV's RANSAC ends the same way: its consensus set is
// |error| < reprojectionError and the pose is refined on it).  One CTA; mask = the inlier flags (in: all ones).
__global__ void __launch_bounds__(kPoseThreads) pose_track_kernel(PoseData d, Params P, double thr, int rounds,
                                                                  unsigned char* mask, int* n_inliers, int classify_first) {
  __shared__ double cred[kPoseWarps], cnt[1];
  __shared__ double cpose[12];
  const int tid = threadIdx.x;
  const long long o0 = d.obs_off[0], o1 = d.obs_off[1];
  Params C = P;                                         // classification: plain pixel error
  C.loss = 0;
  if (classify_first) {                                 // the consensus set of the input pose (a RANSAC hypothesis) at thr
    if (tid < 12) cpose[tid] = d.poses[tid];
    __syncthreads();
    for (long long i = o0 + tid; i < o1; i += kPoseThreads) {
      const double* X = d.points + 3 * (size_t)i;
      const float2 uv = d.obs_uv[i];
      Proj o;
      project(C, cpose, X[0], X[1], X[2], uv.x, uv.y, o);
      mask[i] = (o.valid && o.rho < thr * thr) ? 1 : 0;
    }
    __syncthreads();
  }
  for (int r = 0; r < rounds; ++r) {
    // graduated threshold: thr x 2^(rounds - 2 - r), at least thr (6 rounds: 16, 8, 4, 2, 1, 1) -- the wide rounds pull
    // the pose into the dominant consensus, the narrow ones select it
    const int sh = classify_first ? 0 : rounds - 2 - r;
    const double tr = sh > 0 ? thr * (double)(1 << (sh > 20 ? 20 : sh)) : thr;
    Params Q = P;
    Q.loss = (rounds > 1 && r == rounds - 1) ? 0 : 1;
    Q.huber_a = tr;
    pose_lm_run(d, Q, 0, mask);
    if (tid < 12) cpose[tid] = d.poses[tid];            // written by the same threads at the end of the run
    __syncthreads();
    double c[1] = {0.0};
    for (long long i = o0 + tid; i < o1; i += kPoseThreads) {
      const double* X = d.points + 3 * (size_t)i;
      const float2 uv = d.obs_uv[i];
      Proj o;
      project(C, cpose, X[0], X[1], X[2], uv.x, uv.y, o);
      const bool in = o.valid && o.rho < tr * tr;
      mask[i] = in ? 1 : 0;
      c[0] += in ? 1.0 : 0.0;
    }
    cta_sum<1>(c, cred, cnt);
    const int n = (int)cnt[0];
    if (tid == 0) *n_inliers = n;
    __syncthreads();
    if (n < 6) break;                                   // nothing left to fit: the caller treats it as a failure
  }
}

// ---- sampled consensus (GPU RANSAC) around the pose LM: hypotheses from minimal samples, scored on all observations
__host__ __device__ inline unsigned long long splitmix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// hypothesis j: kSample distinct observations (stateless hash of (seed, j, k); a collision moves on to the next index),
// gathered into problem j of a pose batch that starts from the prior
__global__ void ransac_sample_kernel(int nh, long long n, unsigned long long seed, const double* __restrict__ points,
                                     const float2* __restrict__ uv, const double* __restrict__ prior,
                                     double* hp_pts, float2* hp_uv, double* hp_poses, long long* hp_off, int* hp_idx) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j > nh) return;
  if (j == nh) { hp_off[nh] = (long long)nh * kRansacSample; return; }
  hp_off[j] = (long long)j * kRansacSample;
  long long pick[kRansacSample];
  for (int k = 0; k < kRansacSample; ++k) {
    long long i = (long long)(splitmix64(seed + (unsigned long long)j * kRansacSample + k) % (unsigned long long)n);
    for (bool again = true; again;) {
      again = false;
      for (int q = 0; q < k; ++q) if (pick[q] == i) { i = (i + 1) % n; again = true; }
    }
    pick[k] = i;
    const size_t o = (size_t)j * kRansacSample + k;
    hp_pts[3 * o] = points[3 * i]; hp_pts[3 * o + 1] = points[3 * i + 1]; hp_pts[3 * o + 2] = points[3 * i + 2];
    hp_uv[o] = uv[i];
    hp_idx[o] = (int)i;
  }
  for (int k = 0; k < 12; ++k) hp_poses[12 * (size_t)j + k] = prior[k];
}

// consensus of hypothesis blockIdx.x: observations in front of the camera with |reprojection error| < thr
__global__ void __launch_bounds__(kPoseThreads) ransac_score_kernel(long long n, const double* __restrict__ points,
                                                                    const float2* __restrict__ uv, const double* __restrict__ hp_poses,
                                                                    const b2ba_summary* __restrict__ hp_summ, Params P,
                                                                    double thr, int* score) {
  __shared__ double pose[12];
  __shared__ int red[kPoseWarps];
  const int j = blockIdx.x, tid = threadIdx.x;
  if (tid < 12) pose[tid] = hp_poses[12 * (size_t)j + tid];
  __syncthreads();
  bool fin = true;
  for (int k = 0; k < 12; ++k) fin = fin && isfinite(pose[k]);
  int c = 0;
  if (fin && hp_summ[j].status != B2BA_NUMERICAL_FAILURE) {
    Params C = P;
    C.loss = 0;
    for (long long i = tid; i < n; i += kPoseThreads) {
      Proj o;
      project(C, pose, points[3 * i], points[3 * i + 1], points[3 * i + 2], uv[i].x, uv[i].y, o);
      c += (o.valid && o.rho < thr * thr) ? 1 : 0;
    }
  }
  c = __reduce_add_sync(0xffffffffu, c);
  if ((tid & 31) == 0) red[tid >> 5] = c;
  __syncthreads();
  if (tid == 0) { int t = 0; for (int w = 0; w < kPoseWarps; ++w) t += red[w]; score[j] = t; }
}

// the hypothesis with the largest consensus (ties: the lowest index) becomes pose 0 of the tracking problem
__global__ void ransac_pick_kernel(int nh, const int* __restrict__ score, const double* __restrict__ hp_poses,
                                   double* pose, int* best) {
  __shared__ int bs[256], bi[256];
  const int tid = threadIdx.x;
  int s = -1, idx = 0x7fffffff;
  for (int j = tid; j < nh; j += 256) if (score[j] > s) { s = score[j]; idx = j; }
  bs[tid] = s; bi[tid] = idx;
  __syncthreads();
  for (int m = 128; m > 0; m >>= 1) {
    if (tid < m && (bs[tid + m] > bs[tid] || (bs[tid + m] == bs[tid] && bi[tid + m] < bi[tid]))) { bs[tid] = bs[tid + m]; bi[tid] = bi[tid + m]; }
    __syncthreads();
  }
  if (tid < 12) pose[tid] = hp_poses[12 * (size_t)bi[0] + tid];
  if (tid == 0) { best[0] = bi[0]; best[1] = bs[0]; }
}

}  // namespace

namespace {
// batched two-view linear triangulation + the reference's gates, one thread per match.
// `triangulate_points` (`src/modules/frontend.py:115-148`): cv2.triangulatePoints (DLT: right singular vector of
// the smallest singular value of the 4 x 4 system below), X / w, depth and reprojection error in the SECOND view,
// keep = depth > min_depth && err < max_reproj_err.  The SVD is a one-sided Jacobi (Hestenes) iteration on the
// columns of A with V accumulated, all in registers.
struct TriData {
  int n;
  double P1[12], P2[12];           // K [R | t] of the two views, row-major 3 x 4
  const float2* uv1;
  const float2* uv2;
  double min_depth, max_err;
  double* X;                       // n x 3
  unsigned char* keep;             // n
};

__global__ void __launch_bounds__(128) triangulate_kernel(TriData d) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= d.n) return;
  const float2 a = d.uv1[i], b = d.uv2[i];
  double A[4][4], V[4][4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    A[0][c] = (double)a.x * d.P1[8 + c] - d.P1[c];
    A[1][c] = (double)a.y * d.P1[8 + c] - d.P1[4 + c];
    A[2][c] = (double)b.x * d.P2[8 + c] - d.P2[c];
    A[3][c] = (double)b.y * d.P2[8 + c] - d.P2[4 + c];
#pragma unroll
    for (int r = 0; r < 4; ++r) V[r][c] = r == c ? 1.0 : 0.0;
  }
  for (int sweep = 0; sweep < 12; ++sweep) {
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
        for (int r = 0; r < 4; ++r) { al = fma(A[r][p], A[r][p], al); be = fma(A[r][q], A[r][q], be); ga = fma(A[r][p], A[r][q], ga); }
        if (fabs(ga) > 1e-300 + 1e-17 * sqrt(al * be)) {
          const double zeta = (be - al) / (2.0 * ga);
          const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const double ap = A[r][p], aq = A[r][q], vp = V[r][p], vq = V[r][q];
            A[r][p] = c * ap - s * aq; A[r][q] = s * ap + c * aq;
            V[r][p] = c * vp - s * vq; V[r][q] = s * vp + c * vq;
          }
        }
      }
  }
  double best = 0.0, v[4] = {0, 0, 0, 1};
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    double nrm = 0.0;
#pragma unroll
    for (int r = 0; r < 4; ++r) nrm = fma(A[r][c], A[r][c], nrm);
    if (c == 0 || nrm < best) {
      best = nrm;
#pragma unroll
      for (int r = 0; r < 4; ++r) v[r] = V[r][c];
    }
  }
  const double X0 = v[0] / v[3], X1 = v[1] / v[3], X2 = v[2] / v[3];
  const double h0 = fma(d.P2[0], X0, fma(d.P2[1], X1, fma(d.P2[2], X2, d.P2[3])));
  const double h1 = fma(d.P2[4], X0, fma(d.P2[5], X1, fma(d.P2[6], X2, d.P2[7])));
  const double h2 = fma(d.P2[8], X0, fma(d.P2[9], X1, fma(d.P2[10], X2, d.P2[11])));
  const double e0 = h0 / h2 - (double)b.x, e1 = h1 / h2 - (double)b.y;
  const double err = sqrt(e0 * e0 + e1 * e1);
  d.X[3 * (size_t)i] = X0; d.X[3 * (size_t)i + 1] = X1; d.X[3 * (size_t)i + 2] = X2;
  d.keep[i] = (h2 > d.min_depth && err < d.max_err) ? 1 : 0;    // NaN compares false: degenerate matches are dropped
}
}  // namespace

int launch_triangulate(int n, const double* P1, const double* P2, const float2* uv1, const float2* uv2,
                       double min_depth, double max_err, double* X, unsigned char* keep, cudaStream_t s) {
  if (n <= 0) return 0;
  TriData d;
  d.n = n;
  for (int k = 0; k < 12; ++k) { d.P1[k] = P1[k]; d.P2[k] = P2[k]; }
  d.uv1 = uv1; d.uv2 = uv2; d.min_depth = min_depth; d.max_err = max_err; d.X = X; d.keep = keep;
  triangulate_kernel<<<(n + 127) / 128, 128, 0, s>>>(d);
  return (int)cudaGetLastError();
}

int launch_pose_track(const PoseData& d, const Params& P, double thr, int rounds, unsigned char* mask,
                      int* n_inliers, cudaStream_t s, int classify_first) {
  pose_track_kernel<<<1, kPoseThreads, 0, s>>>(d, P, thr, rounds, mask, n_inliers, classify_first);
  return (int)cudaGetLastError();
}

int launch_pose_ransac(const RansacData& r, const PoseData& track, const Params& P, double thr, cudaStream_t s) {
  ransac_sample_kernel<<<(r.nh + 1 + 127) / 128, 128, 0, s>>>(r.nh, r.n, r.seed, track.points, track.obs_uv, track.poses,
                                                            r.hp_pts, r.hp_uv, r.hp_poses, r.hp_off, r.hp_idx);
  PoseData hd{};
  hd.nb = r.nh; hd.obs_off = r.hp_off; hd.poses = r.hp_poses; hd.points = r.hp_pts; hd.obs_uv = r.hp_uv; hd.summaries = r.hp_summ;
  Params H = P;
  H.loss = 0; H.max_iters = r.hyp_iters; H.ftol = H.gtol = H.ptol = 0.0;
  pose_lm_warp_kernel<<<r.nh, 32, 0, s>>>(hd, H);
  ransac_score_kernel<<<r.nh, kPoseThreads, 0, s>>>(r.n, track.points, track.obs_uv, r.hp_poses, r.hp_summ, P, thr, r.score);
  ransac_pick_kernel<<<1, 256, 0, s>>>(r.nh, r.score, r.hp_poses, track.poses, r.best);
  return (int)cudaGetLastError();
}

int launch_pose_lm(const PoseData& d, const Params& P, cudaStream_t s) {
  if (d.nb <= 0) return 0;
  pose_lm_kernel<<<d.nb, kPoseThreads, 0, s>>>(d, P);
  return (int)cudaGetLastError();
}

}  // namespace b2
