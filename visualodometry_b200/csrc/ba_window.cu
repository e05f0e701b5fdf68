// Window kernel: one CTA = one sliding-window BA problem, whole LM loop on-chip.
//
// Per LM iteration (Spec a0):
//   camera pass   (a1+a2)  camera-major stream -> B_c (6x6 sym), g_c, cost; CTA-wide fixed-order
//                           reduction, no atomics.
//   point pass    (a1-a4)  point-major stream, TPP lanes per landmark: C_p, g_p -> damped 3x3
//                           factor F (C'^-1 = F F^T) -> Z_p = E_k F staged in shared memory; the
//                           Schur complement  sum_p W_p C'^-1 W_p^T = Z Z^T  is accumulated by
//                           FP64 tensor-core MMA (DMMA m8n8k4) with the sum over landmarks as the
//                           K dimension: no atomics, no shuffles, fixed summation order.  An extra
//                           row y_p = F^T g_p makes  Z y = sum E C'^-1 g_p  fall out of the same MMA.
//   solve         (a5)     S = B' - Z Z^T, in-CTA Cholesky + triangular solves.
//   back-subst    (a7)     delta_p, retraction, trial cost, gain ratio, accept/reject, mu update.
#include "ba_window.cuh"

namespace b2 {

__device__ __forceinline__ int packed21(int r, int s) { return r * 6 - r * (r - 1) / 2 + (s - r); }

template <class Cfg>
__global__ void __launch_bounds__(Cfg::NTHREADS, (Cfg::SMEM <= 110 * 1024) ? 2 : 1)
window_lm_kernel(const WinData d, const Params P) {
  constexpr int NTHREADS = Cfg::NTHREADS, NWARPS = Cfg::NWARPS, G = Cfg::G, KS = Cfg::KS;
  constexpr int TPW = Cfg::TPW, TPP = Cfg::TPP, PB = Cfg::PB, LD = Cfg::LD, LDS = Cfg::LDS;
  extern __shared__ __align__(16) double smem[];
  double* pose_cur = smem;
  double* pose_tri = pose_cur + kMaxCam * 12;
  double* Bc = pose_tri + kMaxCam * 12;      // kMaxCam x 21 (upper, packed)
  double* gc = Bc + kMaxCam * 21;            // kMaxCam x 6
  double* Dc = gc + kMaxCam * 6;
  double* dvec = Dc + kMaxCam * 6;           // delta_c (6*nfree)
  double* gred = dvec + 128;
  double* red = gred + 128;                  // NWARPS x 32 reduction scratch
  double* scal = red + Cfg::NRED;            // 16 control scalars
  double* Z = scal + 16;                     // Z tile, later S
  int* fidx = reinterpret_cast<int*>(Z + Cfg::ZS);   // free index or -1
  int* lo = fidx + kMaxCam;                  // # free cameras with index < c  (kMaxCam+1)
  int* iflag = lo + kMaxCam + 1;             // [0] step_bad  [1] accept  [2] stop

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int win = blockIdx.x;
  const int c0 = d.cam_off[win], nc = d.cam_off[win + 1] - c0;
  const int p0 = d.pt_off[win], npw = d.pt_off[win + 1] - p0;

  for (int i = tid; i < nc * 12; i += NTHREADS) pose_cur[i] = d.poses[(size_t)c0 * 12 + i];
  if (tid == 0) {
    int nf = 0;
    for (int c = 0; c < nc; ++c) {
      lo[c] = nf;
      fidx[c] = d.fixed[c0 + c] ? -1 : nf++;
    }
    lo[nc] = nf;
    iflag[0] = iflag[1] = iflag[2] = 0;
  }
  __syncthreads();
  const int nfree = lo[nc];
  const int n = 6 * nfree;              // reduced system size; row n of Z is the y row
  const int nt = (n + 1 + 7) >> 3;      // 8x8 tile rows in use
  const int T = nt * (nt + 1) / 2;
  const int tpw = (T + G - 1) / G;

  // tiles of this warp: group g owns upper tiles [g*tpw, (g+1)*tpw) in row-major upper order
  const int grp = warp % G, ksp = warp / G;
  int tmi[TPW], tni[TPW];
  {
    int t = grp * tpw, mi = 0, rowlen = nt, base = 0;
    while (mi < nt && t >= base + rowlen) { base += rowlen; --rowlen; ++mi; }
    int ni = mi + (t - base);
#pragma unroll
    for (int q = 0; q < TPW; ++q) {
      bool ok = q < tpw && mi < nt;
      tmi[q] = ok ? mi : -1;
      tni[q] = ok ? ni : 0;
      if (++ni >= nt) { ++mi; ni = mi; }
    }
  }

  double mu = P.init_mu, nu = 2.0, cost = 0.0, initial_cost = 0.0;
  int cur = 0, have_lin = 0, it = 0, n_acc = 0, n_rej = 0, status = B2BA_MAX_ITERS, ninv_cur = 0;
  double* const scratch = d.pt_scratch + (size_t)p0 * kPtScratch;
  const int slot = tid / TPP, sub = tid % TPP;
  const int np_round = ((npw + PB - 1) / PB) * PB;
  WinDbg dbgL{nc, npw, n};

  for (; it < P.max_iters; ++it) {
    const double* X = d.pts[cur] + (size_t)p0 * 3;
    double* Xt = d.pts[cur ^ 1] + (size_t)p0 * 3;
    const bool dbg = (d.dbg != nullptr) && it == 0;

    // ------------------------------------------------------------------ camera pass (a1+a2)
    const int fresh_lin = !have_lin;
    if (!have_lin) {
      double cst_tot = 0.0, ninv_tot = 0.0;   // live in thread 27 / 28
      for (int c = 0; c < nc; ++c) {
        const int b = d.cam_ptr[c0 + c], e = d.cam_ptr[c0 + c + 1];
        const double* pose = pose_cur + 12 * c;
        const bool isfree = fidx[c] >= 0;
        double acc[29];
#pragma unroll
        for (int k = 0; k < 29; ++k) acc[k] = 0.0;
        for (int a = b + tid; a < e; a += NTHREADS) {
          const ObsRec rec = d.cm[a];
          const double* Xp = X + 3 * rec.idx;
          Proj o;
          project(P, pose, Xp[0], Xp[1], Xp[2], rec.u, rec.v, o);
          if (!o.valid) { acc[28] += 1.0; continue; }
          acc[27] += 0.5 * o.rho;
          if (isfree) {
            double jc[12];
            jac_cam(P, o, jc);
            int k = 0;
#pragma unroll
            for (int r = 0; r < 6; ++r) {
              const double j0 = o.w * jc[r], j1 = o.w * jc[6 + r];
#pragma unroll
              for (int s = r; s < 6; ++s) acc[k++] += j0 * jc[s] + j1 * jc[6 + s];
              acc[21 + r] += j0 * o.r0 + j1 * o.r1;
            }
          }
        }
#pragma unroll
        for (int k = 0; k < 29; ++k) {
          const double v = warp_sum(acc[k]);
          if (lane == 0) red[warp * 32 + k] = v;
        }
        __syncthreads();
        if (tid < 29) {
          double v = 0.0;
          for (int w = 0; w < NWARPS; ++w) v += red[w * 32 + tid];
          if (tid < 21) Bc[c * 21 + tid] = v;
          else if (tid < 27) gc[c * 6 + tid - 21] = v;
          else if (tid == 27) cst_tot += v;
          else ninv_tot += v;
        }
        __syncthreads();
      }
      if (tid == 27) scal[0] = cst_tot;
      if (tid == 28) scal[1] = ninv_tot;
      for (int i = tid; i < nc * 6; i += NTHREADS) Dc[i] = clampd(Bc[(i / 6) * 21 + packed21(i % 6, i % 6)]);
      __syncthreads();
      cost = scal[0];
      ninv_cur = (int)scal[1];
      if (it == 0) initial_cost = cost;
      have_lin = 1;
    }
    if (dbg) {
      for (int i = tid; i < nc * 36; i += NTHREADS) {
        int c = i / 36, r = (i % 36) / 6, s = i % 6;
        d.dbg[dbgL.off_B() + i] = Bc[c * 21 + packed21(min(r, s), max(r, s))];
      }
      for (int i = tid; i < nc * 6; i += NTHREADS) d.dbg[dbgL.off_gc() + i] = gc[i];
    }

    // ------------------------------------------------------------------ point pass + Schur (a1-a4)
    double acc[TPW][2];
#pragma unroll
    for (int q = 0; q < TPW; ++q) acc[q][0] = acc[q][1] = 0.0;
    // rows above the y row are never written by the point pass: keep them zero
    for (int i = tid; i < (8 * nt - n - 1) * (3 * PB); i += NTHREADS)
      Z[(n + 1 + i / (3 * PB)) * LD + i % (3 * PB)] = 0.0;
    double gmax = 0.0;
    const double imu = 1.0 / mu;

    for (int b0 = 0; b0 < npw; b0 += PB) {
      const int p = b0 + slot;
      const bool act = p < npw;
      int a0 = 0, a1 = 0;
      double X0 = 0, X1 = 0, X2 = 0;
      if (act) {
        a0 = d.pt_ptr[p0 + p]; a1 = d.pt_ptr[p0 + p + 1];
        X0 = X[3 * p]; X1 = X[3 * p + 1]; X2 = X[3 * p + 2];
      }
      // loop 1: C_p, g_p
      double C[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
      for (int a = a0 + sub; a < a1; a += TPP) {
        const ObsRec rec = d.pm[a];
        const double* pose = pose_cur + 12 * rec.idx;
        Proj o;
        project(P, pose, X0, X1, X2, rec.u, rec.v, o);
        double jp[6];
        jac_point(P, pose, o, jp);
        const double w = o.w;
        C[0] += w * (jp[0] * jp[0] + jp[3] * jp[3]); C[1] += w * (jp[0] * jp[1] + jp[3] * jp[4]);
        C[2] += w * (jp[0] * jp[2] + jp[3] * jp[5]); C[3] += w * (jp[1] * jp[1] + jp[4] * jp[4]);
        C[4] += w * (jp[1] * jp[2] + jp[4] * jp[5]); C[5] += w * (jp[2] * jp[2] + jp[5] * jp[5]);
#pragma unroll
        for (int j = 0; j < 3; ++j) g[j] += w * (jp[j] * o.r0 + jp[3 + j] * o.r1);
      }
#pragma unroll
      for (int m = TPP >> 1; m > 0; m >>= 1) {
#pragma unroll
        for (int k = 0; k < 6; ++k) C[k] += __shfl_xor_sync(0xffffffffu, C[k], m);
#pragma unroll
        for (int k = 0; k < 3; ++k) g[k] += __shfl_xor_sync(0xffffffffu, g[k], m);
      }
      const double D0 = clampd(C[0]), D1 = clampd(C[3]), D2 = clampd(C[5]);
      double F[6];
      {
        const double Cd[6] = {C[0] + D0 * imu, C[1], C[2], C[3] + D1 * imu, C[4], C[5] + D2 * imu};
        if (!sym3_inv_factor(Cd, F)) {
#pragma unroll
          for (int k = 0; k < 6; ++k) F[k] = 0.0;
          if (act) iflag[0] = 1;
        }
      }
      gmax = fmax(gmax, fmax(fabs(g[0]), fmax(fabs(g[1]), fabs(g[2]))));
      double* zc = Z + 3 * slot;   // this point's 3 columns
      if (sub == 0) {
        zc[n * LD + 0] = F[0] * g[0];
        zc[n * LD + 1] = F[1] * g[0] + F[3] * g[1];
        zc[n * LD + 2] = F[2] * g[0] + F[4] * g[1] + F[5] * g[2];
        if (act) {
          double* sc = scratch + (size_t)p * kPtScratch;
#pragma unroll
          for (int k = 0; k < 6; ++k) sc[k] = F[k];
          sc[6] = g[0]; sc[7] = g[1]; sc[8] = g[2]; sc[9] = D0; sc[10] = D1; sc[11] = D2;
          if (dbg) {
#pragma unroll
            for (int k = 0; k < 6; ++k) d.dbg[dbgL.off_C() + (size_t)p * 6 + k] = C[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) d.dbg[dbgL.off_gp() + (size_t)p * 3 + k] = g[k];
          }
        }
      }
      // loop 2: Z rows of this point = E_k F at the rows of camera k; zero rows of unseen cameras
      if (a1 == a0) {
        if (sub == 0)
          for (int r = 0; r < n; ++r) { zc[r * LD] = 0.0; zc[r * LD + 1] = 0.0; zc[r * LD + 2] = 0.0; }
      }
      for (int a = a0 + sub; a < a1; a += TPP) {
        const ObsRec rec = d.pm[a];
        const int c = rec.idx, f = fidx[c];
        int zlo = 0;
        if (a > a0) { const int cp = d.pm[a - 1].idx; zlo = lo[cp] + (fidx[cp] >= 0 ? 1 : 0); }
        for (int r = 6 * zlo; r < 6 * lo[c]; ++r) { zc[r * LD] = 0.0; zc[r * LD + 1] = 0.0; zc[r * LD + 2] = 0.0; }
        if (a == a1 - 1)
          for (int r = 6 * (lo[c] + (f >= 0 ? 1 : 0)); r < n; ++r) { zc[r * LD] = 0.0; zc[r * LD + 1] = 0.0; zc[r * LD + 2] = 0.0; }
        if (f < 0) continue;
        const double* pose = pose_cur + 12 * c;
        Proj o;
        project(P, pose, X0, X1, X2, rec.u, rec.v, o);
        double jp[6], jc[12], E[18];
        jac_point(P, pose, o, jp);
        jac_cam(P, o, jc);
        make_E(o.w, jc, jp, E);
        double* zr = zc + 6 * f * LD;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          zr[r * LD + 0] = E[3 * r] * F[0];
          zr[r * LD + 1] = E[3 * r] * F[1] + E[3 * r + 1] * F[3];
          zr[r * LD + 2] = E[3 * r] * F[2] + E[3 * r + 1] * F[4] + E[3 * r + 2] * F[5];
        }
      }
      __syncthreads();
      // Z Z^T accumulate: DMMA over this batch's 3*PB columns (this warp's k-slice)
      {
        const int krow = lane >> 2, kcol = lane & 3;
        const int k_begin = ksp * (Cfg::KSTEPS / KS) * 4, k_end = k_begin + (Cfg::KSTEPS / KS) * 4;
#pragma unroll 2
        for (int k0 = k_begin; k0 < k_end; k0 += 4) {
#pragma unroll
          for (int q = 0; q < TPW; ++q) {
            if (tmi[q] >= 0) {
              const double a = Z[(8 * tmi[q] + krow) * LD + k0 + kcol];
              const double b = Z[(8 * tni[q] + krow) * LD + k0 + kcol];
              dmma884(acc[q][0], acc[q][1], a, b);
            }
          }
        }
      }
      __syncthreads();
    }

    // ------------------------------------------------------------------ S = B' - Z Z^T  (a4)
    double* S = Z;   // the Z tile is dead now; S (n x n, leading dim LDS) aliases it
#pragma unroll
    for (int s = 0; s < KS; ++s) {
      if (ksp == s) {
#pragma unroll
        for (int q = 0; q < TPW; ++q) {
          if (tmi[q] >= 0) {
            double* dst = S + (8 * tmi[q] + (lane >> 2)) * LDS + 8 * tni[q] + 2 * (lane & 3);
            if (s == 0) { dst[0] = acc[q][0]; dst[1] = acc[q][1]; }
            else { dst[0] += acc[q][0]; dst[1] += acc[q][1]; }
          }
        }
      }
      __syncthreads();
    }
    for (int idx = tid; idx < n * n; idx += NTHREADS) {
      const int i = idx / n, j = idx - i * n;
      if (i > j) continue;
      double v = -S[i * LDS + j];
      const int fi = i / 6, fj = j / 6;
      if (fi == fj) {
        // camera index of free slot fi: lo[] is monotone, find c with fidx[c] == fi
        int c = fi;
        while (fidx[c] != fi) ++c;
        const int r = i - 6 * fi, s2 = j - 6 * fj;
        v += Bc[c * 21 + packed21(r, s2)];
        if (r == s2) v += Dc[c * 6 + r] * imu;
      }
      S[i * LDS + j] = v;
      S[j * LDS + i] = v;
    }
    // gred[i] = g_c[i] - (Z y)[i]; (Z y)[i] is column n of the accumulated product (rows i < n)
    if (tid < n) {
      const int fi = tid / 6;
      int c = fi;
      while (fidx[c] != fi) ++c;
      gred[tid] = gc[c * 6 + tid - 6 * fi] - S[tid * LDS + n];
    }
    __syncthreads();
    if (dbg) {
      for (int idx = tid; idx < n * n; idx += NTHREADS) d.dbg[dbgL.off_S() + idx] = S[(idx / n) * LDS + idx % n];
      for (int i = tid; i < n; i += NTHREADS) d.dbg[dbgL.off_gred() + i] = gred[i];
      __syncthreads();
    }

    // ------------------------------------------------------------------ Cholesky + solve (a5)
    for (int j = 0; j < n; ++j) {
      if (tid == 0) {
        double dj = S[j * LDS + j];
        if (!(dj > 0.0) || !isfinite(dj)) { iflag[0] = 1; dj = 1.0; }
        S[j * LDS + j] = sqrt(dj);
      }
      __syncthreads();
      const double inv = 1.0 / S[j * LDS + j];
      for (int i = j + 1 + tid; i < n; i += NTHREADS) S[i * LDS + j] *= inv;
      __syncthreads();
      const int m = n - j - 1;
      for (int idx = tid; idx < m * m; idx += NTHREADS) {
        const int ii = idx / m, kk = idx - ii * m;
        if (kk <= ii) S[(j + 1 + ii) * LDS + j + 1 + kk] -= S[(j + 1 + ii) * LDS + j] * S[(j + 1 + kk) * LDS + j];
      }
      __syncthreads();
    }
    if (tid < n) dvec[tid] = -gred[tid];
    __syncthreads();
    if (warp == 0) {
      for (int i = 0; i < n; ++i) {
        const double xi = dvec[i] / S[i * LDS + i];
        __syncwarp();
        if (lane == 0) dvec[i] = xi;
        for (int k = i + 1 + lane; k < n; k += 32) dvec[k] -= S[k * LDS + i] * xi;
        __syncwarp();
      }
      for (int i = n - 1; i >= 0; --i) {
        const double xi = dvec[i] / S[i * LDS + i];
        __syncwarp();
        if (lane == 0) dvec[i] = xi;
        for (int k = lane; k < i; k += 32) dvec[k] -= S[i * LDS + k] * xi;
        __syncwarp();
      }
    }
    __syncthreads();
    // trial cameras + camera part of the model decrease
    if (tid < nc) {
      const int f = fidx[tid];
      if (f >= 0) retract_pose(pose_cur + 12 * tid, dvec + 6 * f, pose_tri + 12 * tid);
      else
        for (int k = 0; k < 12; ++k) pose_tri[12 * tid + k] = pose_cur[12 * tid + k];
    }
    if (tid == 0) {
      double dDd = 0.0, gd = 0.0, dn2 = 0.0;
      bool fin = true;
      for (int c = 0; c < nc; ++c) {
        const int f = fidx[c];
        if (f < 0) continue;
        for (int r = 0; r < 6; ++r) {
          const double dd = dvec[6 * f + r];
          fin = fin && isfinite(dd);
          dDd += dd * dd * Dc[6 * c + r];
          gd += dd * gc[6 * c + r];
          dn2 += dd * dd;
        }
      }
      if (!fin) iflag[0] = 1;
      scal[2] = dDd; scal[3] = gd; scal[4] = dn2;
    }
    __syncthreads();
    const bool step_ok = iflag[0] == 0;
    if (dbg)
      for (int i = tid; i < n; i += NTHREADS) d.dbg[dbgL.off_dc() + i] = dvec[i];

    // ------------------------------------------------------------------ back-substitution + trial cost (a7)
    double s_tc = 0.0, s_dDd = 0.0, s_gd = 0.0, s_dn2 = 0.0, s_xn2 = 0.0, s_inv = 0.0;
    if (step_ok) {
      for (int p = slot; p < np_round; p += PB) {
        const bool act = p < npw;
        int a0 = 0, a1 = 0;
        double X0 = 0, X1 = 0, X2 = 0;
        if (act) {
          a0 = d.pt_ptr[p0 + p]; a1 = d.pt_ptr[p0 + p + 1];
          X0 = X[3 * p]; X1 = X[3 * p + 1]; X2 = X[3 * p + 2];
        }
        double v[3] = {0, 0, 0};
        for (int a = a0 + sub; a < a1; a += TPP) {
          const ObsRec rec = d.pm[a];
          const int f = fidx[rec.idx];
          if (f < 0) continue;
          const double* pose = pose_cur + 12 * rec.idx;
          Proj o;
          project(P, pose, X0, X1, X2, rec.u, rec.v, o);
          double jp[6], jc[12];
          jac_point(P, pose, o, jp);
          jac_cam(P, o, jc);
          // E^T dc = w J_p^T (J_c dc)
          const double* dc = dvec + 6 * f;
          double t0 = 0.0, t1 = 0.0;
#pragma unroll
          for (int r = 0; r < 6; ++r) { t0 += jc[r] * dc[r]; t1 += jc[6 + r] * dc[r]; }
#pragma unroll
          for (int j = 0; j < 3; ++j) v[j] += o.w * (jp[j] * t0 + jp[3 + j] * t1);
        }
#pragma unroll
        for (int m = TPP >> 1; m > 0; m >>= 1) {
#pragma unroll
          for (int k = 0; k < 3; ++k) v[k] += __shfl_xor_sync(0xffffffffu, v[k], m);
        }
        double Xn0 = X0, Xn1 = X1, Xn2 = X2;
        if (act) {
          const double* sc = scratch + (size_t)p * kPtScratch;
          const double F0 = sc[0], F1 = sc[1], F2 = sc[2], F3 = sc[3], F4 = sc[4], F5 = sc[5];
          const double g0 = sc[6], g1 = sc[7], g2 = sc[8];
          const double t0 = g0 + v[0], t1 = g1 + v[1], t2 = g2 + v[2];
          // delta_p = -F F^T t
          const double u0 = F0 * t0, u1 = F1 * t0 + F3 * t1, u2 = F2 * t0 + F4 * t1 + F5 * t2;
          const double d0 = -(F0 * u0 + F1 * u1 + F2 * u2), d1 = -(F3 * u1 + F4 * u2), d2 = -(F5 * u2);
          Xn0 = X0 + d0; Xn1 = X1 + d1; Xn2 = X2 + d2;
          if (sub == 0) {
            Xt[3 * p] = Xn0; Xt[3 * p + 1] = Xn1; Xt[3 * p + 2] = Xn2;
            s_dDd += d0 * d0 * sc[9] + d1 * d1 * sc[10] + d2 * d2 * sc[11];
            s_gd += d0 * g0 + d1 * g1 + d2 * g2;
            s_dn2 += d0 * d0 + d1 * d1 + d2 * d2;
            s_xn2 += X0 * X0 + X1 * X1 + X2 * X2;
            if (dbg) { double* dp = d.dbg + dbgL.off_dp() + (size_t)p * 3; dp[0] = d0; dp[1] = d1; dp[2] = d2; }
          }
        }
        for (int a = a0 + sub; a < a1; a += TPP) {
          const ObsRec rec = d.pm[a];
          Proj o;
          project(P, pose_tri + 12 * rec.idx, Xn0, Xn1, Xn2, rec.u, rec.v, o);
          if (o.valid) s_tc += 0.5 * o.rho; else s_inv += 1.0;
        }
      }
    }
    {
      double vals[7] = {s_tc, s_dDd, s_gd, s_dn2, s_xn2, s_inv, gmax};
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        const double v = warp_sum(vals[k]);
        if (lane == 0) red[warp * 32 + k] = v;
      }
      double gm = vals[6];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) gm = fmax(gm, __shfl_xor_sync(0xffffffffu, gm, o));
      if (lane == 0) red[warp * 32 + 6] = gm;
    }
    __syncthreads();

    // ------------------------------------------------------------------ LM control (Spec a0)
    if (tid == 0) {
      double tc = 0, dDd = scal[2], gd = scal[3], dn2 = scal[4], xn2 = 0, ninv = 0, gm = 0;
      for (int w = 0; w < NWARPS; ++w) {
        tc += red[w * 32]; dDd += red[w * 32 + 1]; gd += red[w * 32 + 2]; dn2 += red[w * 32 + 3];
        xn2 += red[w * 32 + 4]; ninv += red[w * 32 + 5]; gm = fmax(gm, red[w * 32 + 6]);
      }
      for (int i = 0; i < nc * 6; ++i) if (fidx[i / 6] >= 0) gm = fmax(gm, fabs(gc[i]));
      const double model = 0.5 * (dDd * imu - gd);
      int accept = 0, stop = 0;
      double rho = 0.0;
      if (step_ok && model > 0.0) {
        rho = (cost - tc) / model;
        accept = (isfinite(tc) && rho > kMinRatio) ? 1 : 0;
      }
      // gradient test belongs to the linearisation: stop BEFORE taking this step (not counted)
      if (P.gtol > 0.0 && fresh_lin && gm <= P.gtol) { accept = 0; stop = -1; }
      if (accept) {
        if (P.ftol > 0.0 && fabs(cost - tc) <= P.ftol * cost) stop = B2BA_CONVERGED;
        if (P.ptol > 0.0 && sqrt(dn2) <= P.ptol * (sqrt(xn2) + P.ptol)) stop = B2BA_CONVERGED;
      }
      scal[5] = accept ? tc : cost;
      scal[6] = rho;
      scal[7] = ninv;
      scal[8] = gm;
      if (dbg) {
        double* sc = d.dbg + dbgL.off_scal();
        sc[0] = cost; sc[1] = model; sc[2] = step_ok ? tc : 0.0; sc[3] = step_ok ? 1.0 : 0.0;
      }
      iflag[1] = accept; iflag[2] = stop; iflag[0] = 0;
    }
    __syncthreads();
    const int accept = iflag[1];
    int stop = iflag[2];
    if (stop < 0) { status = B2BA_CONVERGED; break; }
    const double mu_used = mu;
    if (accept) {
      const double rho = scal[6], t = 2.0 * rho - 1.0;
      mu = fmin(kMuMax, mu / fmax(1.0 / 3.0, 1.0 - t * t * t));
      nu = 2.0;
      cost = scal[5];
      ninv_cur = (int)scal[7];
      for (int i = tid; i < nc * 12; i += NTHREADS) pose_cur[i] = pose_tri[i];
      cur ^= 1;
      have_lin = 0;
      ++n_acc;
    } else {
      mu = mu / nu;
      nu *= 2.0;
      ++n_rej;
    }
    if (tid == 0 && d.hist_cost != nullptr && it < kHist) {
      d.hist_cost[(size_t)win * kHist + it] = cost;
      d.hist_mu[(size_t)win * kHist + it] = mu_used;
      d.hist_accept[(size_t)win * kHist + it] = accept;
    }
    if (!stop && !(mu > kMuMin)) stop = B2BA_NUMERICAL_FAILURE;
    __syncthreads();
    if (stop) { status = stop; ++it; break; }
  }

  // ------------------------------------------------------------------ write back
  if (cur == 1) {
    const double* src = d.pts[1] + (size_t)p0 * 3;
    double* dst = d.pts[0] + (size_t)p0 * 3;
    for (int i = tid; i < npw * 3; i += NTHREADS) dst[i] = src[i];
  }
  for (int i = tid; i < nc * 12; i += NTHREADS) d.poses[(size_t)c0 * 12 + i] = pose_cur[i];
  if (tid == 0) {
    b2ba_summary s;
    s.status = status; s.iterations = it; s.accepted = n_acc; s.rejected = n_rej;
    s.n_invalid = ninv_cur; s.pcg_iters = 0;
    s.initial_cost = initial_cost; s.final_cost = cost; s.final_mu = mu;
    d.summaries[win] = s;
  }
}

using CfgSmall = WinCfg<7, 8, 4, 64>;     // n+1 <= 56: up to 9 free cameras (C1 / C4)
using CfgBig = WinCfg<16, 16, 16, 32>;    // n+1 <= 128: up to 21 free cameras (C3)

template <class Cfg>
static int launch(const WinData& d, const Params& p, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(window_lm_kernel<Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)Cfg::SMEM);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  window_lm_kernel<Cfg><<<d.nb, Cfg::NTHREADS, Cfg::SMEM, s>>>(d, p);
  return (int)cudaGetLastError();
}

int launch_window_small(const WinData& d, const Params& p, cudaStream_t s) { return launch<CfgSmall>(d, p, s); }
int launch_window_big(const WinData& d, const Params& p, cudaStream_t s) { return launch<CfgBig>(d, p, s); }

int window_kernel_attrs(int which, int* regs, int* smem_bytes, int* ctas_per_sm) {
  cudaFuncAttributes a;
  cudaError_t e;
  int nb = 0;
  if (which == 0) {
    e = cudaFuncGetAttributes(&a, window_lm_kernel<CfgSmall>);
    if (e == cudaSuccess) {
      cudaFuncSetAttribute(window_lm_kernel<CfgSmall>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgSmall::SMEM);
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, window_lm_kernel<CfgSmall>, CfgSmall::NTHREADS, CfgSmall::SMEM);
    }
    *smem_bytes = (int)CfgSmall::SMEM;
  } else {
    e = cudaFuncGetAttributes(&a, window_lm_kernel<CfgBig>);
    if (e == cudaSuccess) {
      cudaFuncSetAttribute(window_lm_kernel<CfgBig>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgBig::SMEM);
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, window_lm_kernel<CfgBig>, CfgBig::NTHREADS, CfgBig::SMEM);
    }
    *smem_bytes = (int)CfgBig::SMEM;
  }
  if (e != cudaSuccess) return (int)e;
  *regs = a.numRegs;
  *ctas_per_sm = nb;
  return 0;
}

}  // namespace b2
