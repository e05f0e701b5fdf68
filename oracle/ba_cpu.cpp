// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked, imported or executed by the product path
// (visualodometry_b200/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may use it, and only as the checker / the timed CPU baseline.
//
// PARITY UNPINNED: the reference (cteufel13/VisualOdometry) ships NO bundle adjustment
// (SURVEY.md §0), so there is no reference BA output to pin this against.  This file is a
// CPU fp64 restatement of "Spec a0" (SURVEY.md §8): Ceres-style Levenberg-Marquardt with
// point-major Schur elimination.  What IS pinned to the reference: the projection convention
// x_c = R_cw X + t_cw, uv = K x_c / z (src/modules/frontend.py:128-139), the residual
// "projected - observed" in pixels (frontend.py:140), the cheirality gate z > min_depth
// (frontend.py:135, src/config/config.py:31) -- see tests/test_oracle_reference.py, which
// checks these against fixtures generated from the reference's own triangulate_points.
//
// Build: make -C oracle   (g++ -O2 -fopenmp, no third-party linear algebra).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {

struct ba_opts {
  int32_t max_iters;     // LM iterations (one linear solve + one trial evaluation each)
  int32_t loss;          // 0 trivial, 1 Huber
  double loss_scale;     // Huber a (pixels)
  double z_min;          // cheirality gate (config.py:31 min_depth = 1e-3)
  double init_mu;        // initial trust-region radius (1e4)
  double ftol, gtol, ptol;  // 0 disables
  int32_t solver;        // 0 dense Cholesky, 1 PCG (block-Jacobi)
  int32_t pcg_max_iters; // 200
  double pcg_tol;        // 1e-10 on ||r||/||b||
  int32_t threads;       // OpenMP threads (<=0: all)
  int32_t flags;         // bit 0: linearise ONCE per accepted state -- r, w, J_c, J_p of every observation are stored by
                         // linearize() and re-used by the Schur and back-substitution passes (same values, so the
                         // results are bit-identical; this is the form a tuned CPU BA takes and the one timed as the
                         // CPU baseline; 0 keeps the recompute-per-pass form of the first round)
};

struct ba_summary {
  int32_t status;        // 1 max-iters, 2 converged, 3 numerical failure
  int32_t iterations, accepted, rejected, n_invalid, pcg_iters_total;
  double initial_cost, final_cost, final_mu;
  double cost_hist[64], mu_hist[64];   // cost after / mu used at iteration i
  int32_t accept_hist[64];
};
}

namespace {

constexpr double kDiagMin = 1e-6, kDiagMax = 1e32, kMinRatio = 1e-3, kMuMax = 1e16;

struct ObsLin {
  double r[2], w, rho, Jc[12], Jp[6];
  bool valid;
};

// a1: residual, robust weight, analytic 2x6 / 2x3 Jacobians of one observation.
// pose = [R row-major (9) | t (3)], world->camera (frontend.py:134).
inline void obs_eval(const double* pose, const double* X, float uf, float vf, const double* intr,
                     const ba_opts& o, bool want_jac, ObsLin& L) {
  const double* R = pose;
  const double* t = pose + 9;
  double x = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + t[0];
  double y = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + t[1];
  double z = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + t[2];
  L.valid = (z >= o.z_min);
  if (!L.valid) {
    L.r[0] = L.r[1] = 0; L.w = 0; L.rho = 0;
    if (want_jac) { std::fill(L.Jc, L.Jc + 12, 0.0); std::fill(L.Jp, L.Jp + 6, 0.0); }
    return;
  }
  double iz = 1.0 / z;
  double fx = intr[0], fy = intr[1];
  L.r[0] = fx * x * iz + intr[2] - (double)uf;
  L.r[1] = fy * y * iz + intr[3] - (double)vf;
  double s = L.r[0] * L.r[0] + L.r[1] * L.r[1];
  if (o.loss == 1 && s > o.loss_scale * o.loss_scale) {
    double rs = std::sqrt(s);
    L.rho = 2.0 * o.loss_scale * rs - o.loss_scale * o.loss_scale;
    L.w = o.loss_scale / rs;
  } else {
    L.rho = s; L.w = 1.0;
  }
  if (!want_jac) return;
  // d r / d x_c
  double a0 = fx * iz, a2 = -fx * x * iz * iz;
  double b1 = fy * iz, b2 = -fy * y * iz * iz;
  // J_c = dr/dx * [ -[x]x | I ],  -[x]x = [[0,z,-y],[-z,0,x],[y,-x,0]]
  L.Jc[0] = a2 * y;            L.Jc[1] = a0 * z - a2 * x;   L.Jc[2] = -a0 * y;
  L.Jc[3] = a0;                L.Jc[4] = 0;                 L.Jc[5] = a2;
  L.Jc[6] = -b1 * z + b2 * y;  L.Jc[7] = -b2 * x;           L.Jc[8] = b1 * x;
  L.Jc[9] = 0;                 L.Jc[10] = b1;               L.Jc[11] = b2;
  // J_p = dr/dx * R
  for (int j = 0; j < 3; ++j) {
    L.Jp[j] = a0 * R[j] + a2 * R[6 + j];
    L.Jp[3 + j] = b1 * R[3 + j] + b2 * R[6 + j];
  }
}

inline double clampd(double v) { return std::min(std::max(v, kDiagMin), kDiagMax); }

// Exp map (Rodrigues), same branch rule as the engine.
inline void so3_exp(const double* w, double* E) {
  double t2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], a, b;
  if (t2 < 1e-16) { a = 1.0 - t2 / 6.0; b = 0.5 - t2 / 24.0; }
  else { double t = std::sqrt(t2); a = std::sin(t) / t; b = (1.0 - std::cos(t)) / t2; }
  double x = w[0], y = w[1], z = w[2];
  E[0] = 1 + b * (-(y * y + z * z)); E[1] = -a * z + b * x * y;        E[2] = a * y + b * x * z;
  E[3] = a * z + b * x * y;          E[4] = 1 + b * (-(x * x + z * z)); E[5] = -a * x + b * y * z;
  E[6] = -a * y + b * x * z;         E[7] = a * x + b * y * z;          E[8] = 1 + b * (-(x * x + y * y));
}

// pose <- Exp(dw) * pose, t <- Exp(dw) t + dt   (left perturbation, ordering [dw, dt])
inline void retract_pose(const double* pose, const double* d, double* out) {
  double E[9];
  so3_exp(d, E);
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j)
      out[3 * i + j] = E[3 * i] * pose[j] + E[3 * i + 1] * pose[3 + j] + E[3 * i + 2] * pose[6 + j];
    out[9 + i] = E[3 * i] * pose[9] + E[3 * i + 1] * pose[10] + E[3 * i + 2] * pose[11] + d[3 + i];
  }
}

// inverse of a symmetric 3x3 (a3): closed-form adjugate. C = [c00 c01 c02; . c11 c12; . . c22]
inline bool sym3_inv(const double* C, double* I) {
  double c00 = C[0], c01 = C[1], c02 = C[2], c11 = C[3], c12 = C[4], c22 = C[5];
  double m00 = c11 * c22 - c12 * c12, m01 = c02 * c12 - c01 * c22, m02 = c01 * c12 - c02 * c11;
  double det = c00 * m00 + c01 * m01 + c02 * m02;
  if (!(det > 0) || !std::isfinite(det)) return false;
  double id = 1.0 / det;
  I[0] = m00 * id; I[1] = m01 * id; I[2] = m02 * id;
  I[3] = (c00 * c22 - c02 * c02) * id; I[4] = (c01 * c02 - c00 * c12) * id;
  I[5] = (c00 * c11 - c01 * c01) * id;
  return true;
}

}  // namespace

namespace {

// ---------------------------------------------------------------------------------------------
// a8: index builder.  Canonical orders (bit-exact contract with the engine):
//   point-major : observations sorted by (pt, cam, original index)
//   camera-major: observations sorted by (cam, pt, original index)
//   S structure : upper block pairs (fi<=fj) of free-camera indices that share a point, plus all
//                 diagonal blocks, sorted by (fi,fj), CSR by fi.
struct Index {
  int nc = 0, np = 0, nfree = 0;
  int64_t no = 0;
  std::vector<int32_t> pm_perm, pt_ptr, cm_perm, cam_ptr, fidx, blk_ptr, blk_col;
  bool duplicate = false;
};

int build_index(int nc, int np, int64_t no, const int32_t* oc, const int32_t* op,
                const uint8_t* fixed, Index& ix) {
  ix.nc = nc; ix.np = np; ix.no = no;
  for (int64_t i = 0; i < no; ++i)
    if (oc[i] < 0 || oc[i] >= nc || op[i] < 0 || op[i] >= np) return -1;
  ix.pm_perm.resize(no);
  std::iota(ix.pm_perm.begin(), ix.pm_perm.end(), 0);
  std::stable_sort(ix.pm_perm.begin(), ix.pm_perm.end(), [&](int32_t a, int32_t b) {
    return op[a] != op[b] ? op[a] < op[b] : oc[a] < oc[b];
  });
  ix.pt_ptr.assign(np + 1, 0);
  for (int64_t i = 0; i < no; ++i) ix.pt_ptr[op[i] + 1]++;
  for (int p = 0; p < np; ++p) ix.pt_ptr[p + 1] += ix.pt_ptr[p];
  for (int64_t i = 1; i < no; ++i) {
    int32_t a = ix.pm_perm[i - 1], b = ix.pm_perm[i];
    if (op[a] == op[b] && oc[a] == oc[b]) ix.duplicate = true;
  }
  ix.cm_perm = ix.pm_perm;
  std::stable_sort(ix.cm_perm.begin(), ix.cm_perm.end(),
                   [&](int32_t a, int32_t b) { return oc[a] < oc[b]; });
  ix.cam_ptr.assign(nc + 1, 0);
  for (int64_t i = 0; i < no; ++i) ix.cam_ptr[oc[i] + 1]++;
  for (int c = 0; c < nc; ++c) ix.cam_ptr[c + 1] += ix.cam_ptr[c];
  ix.fidx.assign(nc, -1);
  ix.nfree = 0;
  for (int c = 0; c < nc; ++c)
    if (!fixed[c]) ix.fidx[c] = ix.nfree++;
  // S block structure
  std::vector<int64_t> keys;
  keys.reserve(ix.nfree + no * 2);
  for (int f = 0; f < ix.nfree; ++f) keys.push_back((int64_t)f * ix.nfree + f);
  for (int p = 0; p < np; ++p)
    for (int a = ix.pt_ptr[p]; a < ix.pt_ptr[p + 1]; ++a) {
      int fa = ix.fidx[oc[ix.pm_perm[a]]];
      if (fa < 0) continue;
      for (int b = a + 1; b < ix.pt_ptr[p + 1]; ++b) {
        int fb = ix.fidx[oc[ix.pm_perm[b]]];
        if (fb < 0) continue;
        keys.push_back((int64_t)std::min(fa, fb) * ix.nfree + std::max(fa, fb));
      }
    }
  std::sort(keys.begin(), keys.end());
  keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
  ix.blk_ptr.assign(ix.nfree + 1, 0);
  ix.blk_col.resize(keys.size());
  for (size_t i = 0; i < keys.size(); ++i) {
    ix.blk_ptr[keys[i] / ix.nfree + 1]++;
    ix.blk_col[i] = (int32_t)(keys[i] % ix.nfree);
  }
  for (int f = 0; f < ix.nfree; ++f) ix.blk_ptr[f + 1] += ix.blk_ptr[f];
  return 0;
}

inline int find_block(const Index& ix, int fi, int fj) {
  const int32_t* b = ix.blk_col.data() + ix.blk_ptr[fi];
  const int32_t* e = ix.blk_col.data() + ix.blk_ptr[fi + 1];
  return (int)(std::lower_bound(b, e, fj) - ix.blk_col.data());
}

struct Problem {
  int nc, np;
  int64_t no;
  const int32_t *oc, *op;
  const float* uv;
  const uint8_t* fixed;
  const double* intr;
  ba_opts o;
  Index ix;
};

struct Lin {  // one linearisation (a2) + damping + a3
  std::vector<double> B, gc, C, gp, Cinv, Dc, Dp;  // B nc*36 (full sym), C np*6 (upper), Cinv np*6
  std::vector<ObsLin> obs;                         // per observation (original order), only with flags bit 0
  double cost = 0;
  int n_invalid = 0;
};

// the linearisation of observation i: stored (flags bit 0) or recomputed
inline const ObsLin& obs_lin(const Problem& P, const Lin& L, const double* poses, const double* pts, int i, ObsLin& tmp) {
  if (!L.obs.empty()) return L.obs[(size_t)i];
  obs_eval(poses + 12 * P.oc[i], pts + 3 * P.op[i], P.uv[2 * i], P.uv[2 * i + 1], P.intr, P.o, true, tmp);
  return tmp;
}

// a1+a2: cost, B_c, g_c (camera-major pass), C_p, g_p (point-major pass).  Deterministic: each
// block is summed in canonical order by exactly one thread.
void linearize(const Problem& P, const double* poses, const double* pts, Lin& L) {
  const Index& ix = P.ix;
  L.B.assign((size_t)P.nc * 36, 0); L.gc.assign((size_t)P.nc * 6, 0);
  L.C.assign((size_t)P.np * 6, 0);  L.gp.assign((size_t)P.np * 3, 0);
  std::vector<double> ccost(P.nc, 0.0);
  std::vector<int> cinv(P.nc, 0);
  const bool once = (P.o.flags & 1) != 0;
  if (once) {
    L.obs.resize((size_t)P.no);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < P.no; ++i)
      obs_eval(poses + 12 * P.oc[i], pts + 3 * P.op[i], P.uv[2 * i], P.uv[2 * i + 1], P.intr, P.o, true, L.obs[i]);
  } else L.obs.clear();
#pragma omp parallel for schedule(dynamic, 1)
  for (int c = 0; c < P.nc; ++c) {
    double* B = &L.B[(size_t)c * 36];
    double* g = &L.gc[(size_t)c * 6];
    ObsLin tmp;
    for (int a = ix.cam_ptr[c]; a < ix.cam_ptr[c + 1]; ++a) {
      int i = ix.cm_perm[a];
      const ObsLin& o = obs_lin(P, L, poses, pts, i, tmp);
      if (!o.valid) { cinv[c]++; continue; }
      ccost[c] += 0.5 * o.rho;
      for (int r = 0; r < 6; ++r) {
        double j0 = o.w * o.Jc[r], j1 = o.w * o.Jc[6 + r];
        g[r] += j0 * o.r[0] + j1 * o.r[1];
        for (int s = 0; s < 6; ++s) B[6 * r + s] += j0 * o.Jc[s] + j1 * o.Jc[6 + s];
      }
    }
  }
  L.cost = 0; L.n_invalid = 0;
  for (int c = 0; c < P.nc; ++c) { L.cost += ccost[c]; L.n_invalid += cinv[c]; }
#pragma omp parallel for schedule(static)
  for (int p = 0; p < P.np; ++p) {
    double* C = &L.C[(size_t)p * 6];
    double* g = &L.gp[(size_t)p * 3];
    ObsLin tmp;
    for (int a = ix.pt_ptr[p]; a < ix.pt_ptr[p + 1]; ++a) {
      int i = ix.pm_perm[a];
      const ObsLin& o = obs_lin(P, L, poses, pts, i, tmp);
      if (!o.valid) continue;
      const double* J = o.Jp;
      double w = o.w;
      C[0] += w * (J[0] * J[0] + J[3] * J[3]); C[1] += w * (J[0] * J[1] + J[3] * J[4]);
      C[2] += w * (J[0] * J[2] + J[3] * J[5]); C[3] += w * (J[1] * J[1] + J[4] * J[4]);
      C[4] += w * (J[1] * J[2] + J[4] * J[5]); C[5] += w * (J[2] * J[2] + J[5] * J[5]);
      for (int j = 0; j < 3; ++j) g[j] += w * (J[j] * o.r[0] + J[3 + j] * o.r[1]);
    }
  }
}

// damping (Spec a0) + a3
bool damp_and_invert(const Problem& P, Lin& L, double mu) {
  L.Dc.resize((size_t)P.nc * 6); L.Dp.resize((size_t)P.np * 3); L.Cinv.resize((size_t)P.np * 6);
  for (int c = 0; c < P.nc; ++c)
    for (int r = 0; r < 6; ++r) L.Dc[6 * c + r] = clampd(L.B[(size_t)c * 36 + 7 * r]);
  bool ok = true;
#pragma omp parallel for schedule(static) reduction(&& : ok)
  for (int p = 0; p < P.np; ++p) {
    const double* C = &L.C[(size_t)p * 6];
    double* D = &L.Dp[(size_t)p * 3];
    D[0] = clampd(C[0]); D[1] = clampd(C[3]); D[2] = clampd(C[5]);
    double Cd[6] = {C[0] + D[0] / mu, C[1], C[2], C[3] + D[1] / mu, C[4], C[5] + D[2] / mu};
    ok = sym3_inv(Cd, &L.Cinv[(size_t)p * 6]) && ok;
  }
  return ok;
}

}  // namespace

namespace {

// E_k = w J_c^T J_p (6x3, row-major) for one observation
inline void make_E(const ObsLin& o, double* E) {
  for (int r = 0; r < 6; ++r)
    for (int j = 0; j < 3; ++j)
      E[3 * r + j] = o.w * (o.Jc[r] * o.Jp[j] + o.Jc[6 + r] * o.Jp[3 + j]);
}
inline void mul_E_sym3(const double* E, const double* I, double* Y) {  // Y = E * Cinv (sym upper)
  for (int r = 0; r < 6; ++r) {
    double e0 = E[3 * r], e1 = E[3 * r + 1], e2 = E[3 * r + 2];
    Y[3 * r] = e0 * I[0] + e1 * I[1] + e2 * I[2];
    Y[3 * r + 1] = e0 * I[1] + e1 * I[3] + e2 * I[4];
    Y[3 * r + 2] = e0 * I[2] + e1 * I[4] + e2 * I[5];
  }
}

// a4: reduced camera system in block-sparse upper form.  Sv: 36 doubles per block of ix.blk_col
// (row-major 6x6); gred: 6*nfree.   S = B' - sum_p sum_{k,k'} E_k C'^-1 E_k'^T.
void schur(const Problem& P, const double* poses, const double* pts, const Lin& L, double mu,
           std::vector<double>& Sv, std::vector<double>& gred) {
  const Index& ix = P.ix;
  Sv.assign(ix.blk_col.size() * 36, 0.0);
  gred.assign((size_t)ix.nfree * 6, 0.0);
  for (int c = 0; c < P.nc; ++c) {
    int f = ix.fidx[c];
    if (f < 0) continue;
    double* D = &Sv[(size_t)find_block(ix, f, f) * 36];
    for (int r = 0; r < 6; ++r) {
      for (int s = 0; s < 6; ++s) D[6 * r + s] = L.B[(size_t)c * 36 + 6 * r + s];
      D[7 * r] += L.Dc[6 * c + r] / mu;
      gred[6 * f + r] = L.gc[6 * c + r];
    }
  }
  int nthr = 1;
#ifdef _OPENMP
  nthr = omp_in_parallel() ? 1 : omp_get_max_threads();
#endif
  // thread-private accumulators (no atomics) when they are small; else atomics on the shared S
  const bool priv = nthr == 1 || Sv.size() * (size_t)nthr <= ((size_t)8 << 20);
  std::vector<std::vector<double>> Sp(priv && nthr > 1 ? nthr : 0), gp_(priv && nthr > 1 ? nthr : 0);
#pragma omp parallel num_threads(nthr)
  {
    int tid = 0;
#ifdef _OPENMP
    tid = omp_get_thread_num();
#endif
    double* Sacc = Sv.data();
    double* gacc = gred.data();
    if (priv && nthr > 1) {
      Sp[tid].assign(Sv.size(), 0.0); gp_[tid].assign(gred.size(), 0.0);
      Sacc = Sp[tid].data(); gacc = gp_[tid].data();
    }
    std::vector<double> Es, Ys;
    std::vector<int> fs;
#pragma omp for schedule(static)
    for (int p = 0; p < P.np; ++p) {
      int a0 = ix.pt_ptr[p], a1 = ix.pt_ptr[p + 1];
      Es.clear(); Ys.clear(); fs.clear();
      ObsLin tmp;
      for (int a = a0; a < a1; ++a) {
        int i = ix.pm_perm[a];
        int f = ix.fidx[P.oc[i]];
        if (f < 0) continue;
        const ObsLin& o = obs_lin(P, L, poses, pts, i, tmp);
        if (!o.valid) continue;
        size_t k = fs.size();
        fs.push_back(f);
        Es.resize(18 * (k + 1)); Ys.resize(18 * (k + 1));
        make_E(o, &Es[18 * k]);
        mul_E_sym3(&Es[18 * k], &L.Cinv[(size_t)p * 6], &Ys[18 * k]);
      }
      const double* gp = &L.gp[(size_t)p * 3];
      for (size_t k = 0; k < fs.size(); ++k) {
        const double* Y = &Ys[18 * k];
        for (int r = 0; r < 6; ++r) {
          double v = Y[3 * r] * gp[0] + Y[3 * r + 1] * gp[1] + Y[3 * r + 2] * gp[2];
          if (priv) gacc[6 * fs[k] + r] -= v;
          else {
#pragma omp atomic
            gacc[6 * fs[k] + r] -= v;
          }
        }
        for (size_t k2 = k; k2 < fs.size(); ++k2) {  // fs is increasing (cam order, no duplicates)
          const double* E2 = &Es[18 * k2];
          double* Sb = &Sacc[(size_t)find_block(ix, fs[k], fs[k2]) * 36];
          for (int r = 0; r < 6; ++r)
            for (int s = 0; s < 6; ++s) {
              double v = Y[3 * r] * E2[3 * s] + Y[3 * r + 1] * E2[3 * s + 1] + Y[3 * r + 2] * E2[3 * s + 2];
              if (priv) Sb[6 * r + s] -= v;
              else {
#pragma omp atomic
                Sb[6 * r + s] -= v;
              }
            }
        }
      }
    }
  }
  for (size_t t = 0; t < Sp.size(); ++t) {  // fixed-order reduction of the private copies
    for (size_t i = 0; i < Sv.size(); ++i) Sv[i] += Sp[t][i];
    for (size_t i = 0; i < gred.size(); ++i) gred[i] += gp_[t][i];
  }
}

// a5: dense Cholesky solve of S x = rhs (S expanded from the upper blocks). false if not SPD.
bool solve_dense(const Index& ix, const std::vector<double>& Sv, const std::vector<double>& rhs,
                 std::vector<double>& x) {
  int n = 6 * ix.nfree;
  std::vector<double> A((size_t)n * n, 0.0);
  for (int fi = 0; fi < ix.nfree; ++fi)
    for (int b = ix.blk_ptr[fi]; b < ix.blk_ptr[fi + 1]; ++b) {
      int fj = ix.blk_col[b];
      for (int r = 0; r < 6; ++r)
        for (int s = 0; s < 6; ++s) {
          double v = Sv[(size_t)b * 36 + 6 * r + s];
          if (fi == fj) { if (s >= r) { A[(size_t)(6 * fi + r) * n + 6 * fj + s] = v; A[(size_t)(6 * fj + s) * n + 6 * fi + r] = v; } }
          else { A[(size_t)(6 * fi + r) * n + 6 * fj + s] = v; A[(size_t)(6 * fj + s) * n + 6 * fi + r] = v; }
        }
    }
  // lower Cholesky, row-oriented
  for (int j = 0; j < n; ++j) {
    double d = A[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) d -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
    if (!(d > 0) || !std::isfinite(d)) return false;
    d = std::sqrt(d);
    A[(size_t)j * n + j] = d;
#pragma omp parallel for schedule(static) if (n > 512)
    for (int i = j + 1; i < n; ++i) {
      double v = A[(size_t)i * n + j];
      for (int k = 0; k < j; ++k) v -= A[(size_t)i * n + k] * A[(size_t)j * n + k];
      A[(size_t)i * n + j] = v / d;
    }
  }
  x = rhs;
  for (int i = 0; i < n; ++i) {
    double v = x[i];
    for (int k = 0; k < i; ++k) v -= A[(size_t)i * n + k] * x[k];
    x[i] = v / A[(size_t)i * n + i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double v = x[i];
    for (int k = i + 1; k < n; ++k) v -= A[(size_t)k * n + i] * x[k];
    x[i] = v / A[(size_t)i * n + i];
  }
  return true;
}

}  // namespace

namespace {

// y = S x with S symmetric, stored as upper blocks
void bsr_sym_matvec(const Index& ix, const std::vector<double>& Sv, const double* x, double* y) {
  int n = 6 * ix.nfree;
  std::fill(y, y + n, 0.0);
  for (int fi = 0; fi < ix.nfree; ++fi)
    for (int b = ix.blk_ptr[fi]; b < ix.blk_ptr[fi + 1]; ++b) {
      int fj = ix.blk_col[b];
      const double* M = &Sv[(size_t)b * 36];
      if (fi == fj) {
        for (int r = 0; r < 6; ++r)
          for (int s = 0; s < 6; ++s) y[6 * fi + r] += M[6 * std::min(r, s) + std::max(r, s)] * x[6 * fi + s];
      } else {
        for (int r = 0; r < 6; ++r)
          for (int s = 0; s < 6; ++s) {
            y[6 * fi + r] += M[6 * r + s] * x[6 * fj + s];
            y[6 * fj + s] += M[6 * r + s] * x[6 * fi + r];
          }
      }
    }
}

// 6x6 SPD inverse via Cholesky (block-Jacobi preconditioner); upper triangle of M is read
bool inv6(const double* M, double* Inv) {
  double Lm[36] = {0};
  for (int j = 0; j < 6; ++j) {
    double d = M[6 * j + j];
    for (int k = 0; k < j; ++k) d -= Lm[6 * j + k] * Lm[6 * j + k];
    if (!(d > 0) || !std::isfinite(d)) return false;
    d = std::sqrt(d); Lm[6 * j + j] = d;
    for (int i = j + 1; i < 6; ++i) {
      double v = M[6 * j + i];
      for (int k = 0; k < j; ++k) v -= Lm[6 * i + k] * Lm[6 * j + k];
      Lm[6 * i + j] = v / d;
    }
  }
  for (int c = 0; c < 6; ++c) {
    double e[6] = {0}; e[c] = 1;
    for (int i = 0; i < 6; ++i) { double v = e[i]; for (int k = 0; k < i; ++k) v -= Lm[6 * i + k] * e[k]; e[i] = v / Lm[7 * i]; }
    for (int i = 5; i >= 0; --i) { double v = e[i]; for (int k = i + 1; k < 6; ++k) v -= Lm[6 * k + i] * e[k]; e[i] = v / Lm[7 * i]; }
    for (int i = 0; i < 6; ++i) Inv[6 * i + c] = e[i];
  }
  return true;
}

// a6: PCG, block-Jacobi preconditioner, x0 = 0, stop at ||r|| <= tol ||b|| or max iterations.
bool solve_pcg(const Index& ix, const std::vector<double>& Sv, const std::vector<double>& rhs,
               std::vector<double>& x, int max_it, double tol, int& iters) {
  int n = 6 * ix.nfree;
  std::vector<double> Minv((size_t)ix.nfree * 36), r(rhs), z(n), p(n), q(n);
  for (int f = 0; f < ix.nfree; ++f)
    if (!inv6(&Sv[(size_t)find_block(ix, f, f) * 36], &Minv[(size_t)f * 36])) return false;
  auto precond = [&](const double* in, double* out) {
    for (int f = 0; f < ix.nfree; ++f)
      for (int r_ = 0; r_ < 6; ++r_) {
        double v = 0;
        for (int s = 0; s < 6; ++s) v += Minv[(size_t)f * 36 + 6 * r_ + s] * in[6 * f + s];
        out[6 * f + r_] = v;
      }
  };
  auto dot = [&](const double* a, const double* b) { double v = 0; for (int i = 0; i < n; ++i) v += a[i] * b[i]; return v; };
  x.assign(n, 0.0);
  double bnorm = std::sqrt(dot(rhs.data(), rhs.data()));
  iters = 0;
  if (bnorm == 0) return true;
  precond(r.data(), z.data());
  p = z;
  double rz = dot(r.data(), z.data());
  for (int it = 0; it < max_it; ++it) {
    if (std::sqrt(dot(r.data(), r.data())) <= tol * bnorm) break;
    bsr_sym_matvec(ix, Sv, p.data(), q.data());
    double pq = dot(p.data(), q.data());
    if (!(pq > 0) || !std::isfinite(pq)) return false;
    double alpha = rz / pq;
    for (int i = 0; i < n; ++i) { x[i] += alpha * p[i]; r[i] -= alpha * q[i]; }
    precond(r.data(), z.data());
    double rz2 = dot(r.data(), z.data());
    double beta = rz2 / rz;
    rz = rz2;
    for (int i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
    iters = it + 1;
  }
  return true;
}

// a7 (first half): landmark back-substitution; returns delta_p and accumulates the
// model-decrease terms dDd = delta^T D delta (points part) and gd = g^T delta (points part).
void back_substitute(const Problem& P, const double* poses, const double* pts, const Lin& L,
                     const std::vector<double>& dc, std::vector<double>& dp, double& dDd, double& gd) {
  const Index& ix = P.ix;
  dp.assign((size_t)P.np * 3, 0.0);
  double s1 = 0, s2 = 0;
#pragma omp parallel for schedule(static) reduction(+ : s1, s2)
  for (int p = 0; p < P.np; ++p) {
    double v[3] = {L.gp[3 * (size_t)p], L.gp[3 * (size_t)p + 1], L.gp[3 * (size_t)p + 2]};
    ObsLin tmp;
    double E[18];
    for (int a = ix.pt_ptr[p]; a < ix.pt_ptr[p + 1]; ++a) {
      int i = ix.pm_perm[a];
      int f = ix.fidx[P.oc[i]];
      if (f < 0) continue;
      const ObsLin& o = obs_lin(P, L, poses, pts, i, tmp);
      if (!o.valid) continue;
      make_E(o, E);
      for (int r = 0; r < 6; ++r)
        for (int j = 0; j < 3; ++j) v[j] += E[3 * r + j] * dc[6 * f + r];
    }
    const double* I = &L.Cinv[(size_t)p * 6];
    double d[3] = {-(I[0] * v[0] + I[1] * v[1] + I[2] * v[2]), -(I[1] * v[0] + I[3] * v[1] + I[4] * v[2]),
                   -(I[2] * v[0] + I[4] * v[1] + I[5] * v[2])};
    for (int j = 0; j < 3; ++j) {
      dp[3 * (size_t)p + j] = d[j];
      s1 += d[j] * d[j] * L.Dp[3 * (size_t)p + j];
      s2 += d[j] * L.gp[3 * (size_t)p + j];
    }
  }
  dDd = s1; gd = s2;
}

double total_cost(const Problem& P, const double* poses, const double* pts, int* n_invalid) {
  double c = 0;
  int bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : c, bad)
  for (int64_t i = 0; i < P.no; ++i) {
    ObsLin o;
    obs_eval(poses + 12 * P.oc[i], pts + 3 * P.op[i], P.uv[2 * i], P.uv[2 * i + 1], P.intr, P.o, false, o);
    if (o.valid) c += 0.5 * o.rho; else bad++;
  }
  if (n_invalid) *n_invalid = bad;
  return c;
}

}  // namespace

namespace {

struct Step {  // everything one LM iteration produces before the accept test
  std::vector<double> Sv, gred, dc, dp;
  double model_decrease = 0;
  bool ok = false;
  int pcg_iters = 0;
};

void compute_step(const Problem& P, const double* poses, const double* pts, Lin& L, double mu, Step& st) {
  const Index& ix = P.ix;
  st.ok = damp_and_invert(P, L, mu);
  if (!st.ok) return;
  schur(P, poses, pts, L, mu, st.Sv, st.gred);
  std::vector<double> rhs(st.gred.size());
  for (size_t i = 0; i < rhs.size(); ++i) rhs[i] = -st.gred[i];
  if (P.o.solver == 1) st.ok = solve_pcg(ix, st.Sv, rhs, st.dc, P.o.pcg_max_iters, P.o.pcg_tol, st.pcg_iters);
  else st.ok = solve_dense(ix, st.Sv, rhs, st.dc);
  if (!st.ok) return;
  double dDd = 0, gd = 0;
  back_substitute(P, poses, pts, L, st.dc, st.dp, dDd, gd);
  for (int c = 0; c < P.nc; ++c) {
    int f = ix.fidx[c];
    if (f < 0) continue;
    for (int r = 0; r < 6; ++r) {
      double d = st.dc[6 * f + r];
      dDd += d * d * L.Dc[6 * c + r];
      gd += d * L.gc[6 * c + r];
    }
  }
  st.model_decrease = 0.5 * (dDd / mu - gd);
  for (double v : st.dc) if (!std::isfinite(v)) st.ok = false;
}

void apply_step(const Problem& P, const double* poses, const double* pts, const Step& st,
                double* nposes, double* npts) {
  for (int c = 0; c < P.nc; ++c) {
    int f = P.ix.fidx[c];
    if (f < 0) std::memcpy(nposes + 12 * c, poses + 12 * c, 12 * sizeof(double));
    else retract_pose(poses + 12 * c, &st.dc[6 * f], nposes + 12 * c);
  }
  for (size_t i = 0; i < (size_t)P.np * 3; ++i) npts[i] = pts[i] + st.dp[i];
}

int setup(Problem& P, int nc, int np, int64_t no, const int32_t* oc, const int32_t* op, const float* uv,
          const uint8_t* fixed, const double* intr, const ba_opts* o) {
  P.nc = nc; P.np = np; P.no = no; P.oc = oc; P.op = op; P.uv = uv; P.fixed = fixed; P.intr = intr; P.o = *o;
#ifdef _OPENMP
  omp_set_num_threads(o->threads > 0 ? o->threads : omp_get_num_procs());
#endif
  if (build_index(nc, np, no, oc, op, fixed, P.ix) != 0) return -1;
  if (P.ix.duplicate) return -2;
  return 0;
}

int lm_solve(Problem& P, double* poses, double* pts, ba_summary* sm) {
  std::memset(sm, 0, sizeof(*sm));
  double mu = P.o.init_mu, nu = 2.0;
  std::vector<double> tposes((size_t)P.nc * 12), tpts((size_t)P.np * 3);
  Lin L;
  Step st;
  int it = 0;
  sm->status = 1;
  bool have_lin = false;
  double cost = 0;
  int ninv_cur = 0;   // observations behind the cheirality gate at the current state
  for (; it < P.o.max_iters; ++it) {
    if (!have_lin) {  // re-linearise only after an accepted step (same state => same values)
      linearize(P, poses, pts, L);
      have_lin = true;
      cost = L.cost;
      ninv_cur = L.n_invalid;
      if (it == 0) sm->initial_cost = cost;
      if (P.o.gtol > 0) {
        double gmax = 0;
        for (double v : L.gc) gmax = std::max(gmax, std::fabs(v));
        for (double v : L.gp) gmax = std::max(gmax, std::fabs(v));
        if (gmax <= P.o.gtol) { sm->status = 2; break; }
      }
    }
    compute_step(P, poses, pts, L, mu, st);
    sm->pcg_iters_total += st.pcg_iters;
    bool accept = false;
    double tcost = cost, rho = 0;
    if (st.ok && st.model_decrease > 0) {
      apply_step(P, poses, pts, st, tposes.data(), tpts.data());
      int ninv_trial = 0;
      tcost = total_cost(P, tposes.data(), tpts.data(), &ninv_trial);
      rho = (cost - tcost) / st.model_decrease;
      // An observation behind the cheirality gate costs 0, so a step that pushes MORE observations behind it would be
      // rewarded for discarding them: such a step is rejected like a failed evaluation and the radius shrinks (the
      // behaviour of a Ceres cost functor that returns false for z < z_min).  n_invalid never grows.
      accept = std::isfinite(tcost) && rho > kMinRatio && ninv_trial <= ninv_cur;
      if (accept) ninv_cur = ninv_trial;
    }
    if (it < 64) sm->mu_hist[it] = mu;
    bool converged = false;
    if (accept) {
      if (P.o.ftol > 0 && std::fabs(cost - tcost) <= P.o.ftol * cost) converged = true;
      if (P.o.ptol > 0) {
        double dn = 0, xn = 0;
        for (double v : st.dc) dn += v * v;
        for (double v : st.dp) dn += v * v;
        for (size_t i = 0; i < (size_t)P.np * 3; ++i) xn += pts[i] * pts[i];
        if (std::sqrt(dn) <= P.o.ptol * (std::sqrt(xn) + P.o.ptol)) converged = true;
      }
      std::memcpy(poses, tposes.data(), tposes.size() * sizeof(double));
      std::memcpy(pts, tpts.data(), tpts.size() * sizeof(double));
      cost = tcost;
      double t = 2.0 * rho - 1.0;
      mu = std::min(kMuMax, mu / std::max(1.0 / 3.0, 1.0 - t * t * t));
      nu = 2.0;
      have_lin = false;
      sm->accepted++;
    } else {
      mu = mu / nu;
      nu *= 2.0;
      sm->rejected++;
    }
    if (it < 64) { sm->cost_hist[it] = cost; sm->accept_hist[it] = accept ? 1 : 0; }
    if (converged) { sm->status = 2; ++it; break; }
    if (!(mu > 1e-32)) { sm->status = 3; ++it; break; }
  }
  sm->iterations = it;
  sm->final_cost = cost;
  sm->final_mu = mu;
  total_cost(P, poses, pts, &sm->n_invalid);
  return sm->status;
}

}  // namespace

extern "C" {

// Full LM solve of one problem; poses (nc*12) / points (np*3) are updated in place.
int ba_cpu_solve(int nc, int np, int64_t no, double* poses, double* points, const int32_t* obs_cam,
                 const int32_t* obs_pt, const float* obs_uv, const uint8_t* fixed, const double* intr,
                 const ba_opts* o, ba_summary* sm) {
  Problem P;
  int rc = setup(P, nc, np, no, obs_cam, obs_pt, obs_uv, fixed, intr, o);
  if (rc) return rc;
  return lm_solve(P, poses, points, sm);
}

// B independent windows (C4), CSR-packed, indices window-local; OpenMP across windows.
int ba_cpu_solve_batch(int nb, const int32_t* cam_off, const int32_t* pt_off, const int64_t* obs_off,
                       double* poses, double* points, const int32_t* obs_cam, const int32_t* obs_pt,
                       const float* obs_uv, const uint8_t* fixed, const double* intr, const ba_opts* o,
                       ba_summary* sms) {
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = o->threads > 0 ? o->threads : omp_get_num_procs();
  omp_set_max_active_levels(1);
#endif
  int bad = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) reduction(+ : bad)
  for (int b = 0; b < nb; ++b) {
    Problem P;
    ba_opts ob = *o;
    ob.threads = 1;
    int64_t o0 = obs_off[b];
    P.nc = cam_off[b + 1] - cam_off[b]; P.np = pt_off[b + 1] - pt_off[b]; P.no = obs_off[b + 1] - o0;
    P.oc = obs_cam + o0; P.op = obs_pt + o0; P.uv = obs_uv + 2 * o0; P.fixed = fixed + cam_off[b];
    P.intr = intr; P.o = ob;
    if (build_index(P.nc, P.np, P.no, P.oc, P.op, P.fixed, P.ix) != 0 || P.ix.duplicate) { bad++; continue; }
    lm_solve(P, poses + 12 * (size_t)cam_off[b], points + 3 * (size_t)pt_off[b], &sms[b]);
  }
  return bad ? -1 : 0;
}

// a8 dump: index arrays (each may be NULL).  Returns the number of S blocks, <0 on error.
int64_t ba_cpu_index(int nc, int np, int64_t no, const int32_t* obs_cam, const int32_t* obs_pt,
                     const uint8_t* fixed, int32_t* pm_perm, int32_t* pt_ptr, int32_t* cm_perm,
                     int32_t* cam_ptr, int32_t* fidx, int32_t* blk_ptr, int32_t* blk_col) {
  Index ix;
  if (build_index(nc, np, no, obs_cam, obs_pt, fixed, ix) != 0) return -1;
  if (ix.duplicate) return -2;
  auto cp = [](int32_t* dst, const std::vector<int32_t>& v) { if (dst) std::memcpy(dst, v.data(), v.size() * 4); };
  cp(pm_perm, ix.pm_perm); cp(pt_ptr, ix.pt_ptr); cp(cm_perm, ix.cm_perm); cp(cam_ptr, ix.cam_ptr);
  cp(fidx, ix.fidx); cp(blk_ptr, ix.blk_ptr); cp(blk_col, ix.blk_col);
  return (int64_t)ix.blk_col.size();
}

// Per-stage intermediates of ONE LM iteration at radius mu (all outputs optional / may be NULL):
//  r (no*2), w (no), Jc (no*12), Jp (no*6) in ORIGINAL observation order; B (nc*36), gc (nc*6),
//  C (np*6 upper), gp (np*3); S dense (6nfree)^2 row-major; gred, dc (6nfree); dp (np*3);
//  scalars[0]=cost, [1]=model_decrease, [2]=trial cost, [3]=step ok; tposes, tpoints = trial state.
int ba_cpu_stage(int nc, int np, int64_t no, const double* poses, const double* points,
                 const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv, const uint8_t* fixed,
                 const double* intr, const ba_opts* o, double mu, double* r, double* w, double* Jc,
                 double* Jp, double* B, double* gc, double* C, double* gp, double* S, double* gred,
                 double* dc, double* dp, double* scalars, double* tposes, double* tpoints) {
  Problem P;
  int rc = setup(P, nc, np, no, obs_cam, obs_pt, obs_uv, fixed, intr, o);
  if (rc) return rc;
  for (int64_t i = 0; i < no; ++i) {
    ObsLin ol;
    obs_eval(poses + 12 * obs_cam[i], points + 3 * obs_pt[i], obs_uv[2 * i], obs_uv[2 * i + 1], intr, P.o, true, ol);
    if (r) { r[2 * i] = ol.r[0]; r[2 * i + 1] = ol.r[1]; }
    if (w) w[i] = ol.w;
    if (Jc) std::memcpy(Jc + 12 * i, ol.Jc, 12 * sizeof(double));
    if (Jp) std::memcpy(Jp + 6 * i, ol.Jp, 6 * sizeof(double));
  }
  Lin L;
  linearize(P, poses, points, L);
  Step st;
  compute_step(P, poses, points, L, mu, st);
  auto cp = [](double* dst, const std::vector<double>& v) { if (dst && !v.empty()) std::memcpy(dst, v.data(), v.size() * 8); };
  cp(B, L.B); cp(gc, L.gc); cp(C, L.C); cp(gp, L.gp); cp(gred, st.gred); cp(dc, st.dc); cp(dp, st.dp);
  int n = 6 * P.ix.nfree;
  if (S && !st.Sv.empty()) {
    std::fill(S, S + (size_t)n * n, 0.0);
    for (int fi = 0; fi < P.ix.nfree; ++fi)
      for (int b = P.ix.blk_ptr[fi]; b < P.ix.blk_ptr[fi + 1]; ++b) {
        int fj = P.ix.blk_col[b];
        for (int a = 0; a < 6; ++a)
          for (int c = 0; c < 6; ++c) {
            if (fi == fj && c < a) continue;
            double v = st.Sv[(size_t)b * 36 + 6 * a + c];
            S[(size_t)(6 * fi + a) * n + 6 * fj + c] = v;
            S[(size_t)(6 * fj + c) * n + 6 * fi + a] = v;
          }
      }
  }
  double tc = 0;
  if (st.ok) {
    std::vector<double> tp((size_t)nc * 12), tx((size_t)np * 3);
    apply_step(P, poses, points, st, tp.data(), tx.data());
    tc = total_cost(P, tp.data(), tx.data(), nullptr);
    cp(tposes, tp); cp(tpoints, tx);
  }
  if (scalars) { scalars[0] = L.cost; scalars[1] = st.model_decrease; scalars[2] = tc; scalars[3] = st.ok ? 1 : 0; }
  return 0;
}

int ba_cpu_num_procs(void) {
#ifdef _OPENMP
  return omp_get_num_procs();
#else
  return 1;
#endif
}
}
