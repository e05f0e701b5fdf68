// Window kernel: one CTA = one sliding-window BA problem, whole LM loop on-chip.
//
// Per LM iteration (Spec a0):
//   camera pass   (a1+a2)  camera-major stream -> B_c (6x6 sym), g_c, cost; CTA-wide fixed-order
//                           reduction, no atomics.
//   point pass    (a1-a4)  point-major stream, TPP lanes per landmark: C_p, g_p -> damped 3x3
//                           factor F (C'^-1 = F F^T) -> Z_p = E_k F staged in shared memory
//                           (transposed: one contiguous column block per landmark); the Schur
//                           complement  sum_p W_p C'^-1 W_p^T = Z Z^T  is accumulated by FP64
//                           tensor-core MMA (DMMA m8n8k4) with the sum over landmarks as the K
//                           dimension: no atomics, no shuffles, fixed summation order.  An extra
//                           row y_p = F^T g_p makes  Z y = sum E C'^-1 g_p  fall out of the same MMA.
//                           The tile schedule is compile-time per warp group (fragments of a tile
//                           row are loaded once and reused across its tiles).
//   solve         (a5)     S = B' - Z Z^T, in-CTA Cholesky + triangular solves.
//   back-subst    (a7)     delta_p, retraction, trial cost, gain ratio, accept/reject, mu update.
#include "ba_window.cuh"

#include <cooperative_groups.h>

#include <cstdlib>
#include <mutex>
#include <utility>

namespace b2 {

namespace cg = cooperative_groups;

// experiment switches of the window kernel (defaults = the shipped configuration)
#ifndef WX_STASH
#define WX_STASH 0      // loop 2 reads w J_p / J_c parked by loop 1 in the Z slot instead of re-projecting
#endif
#ifndef WX_RECPIPE_PT
#define WX_RECPIPE_PT 1 // point pass: first two records of the next batch loaded into registers during the DMMA
#endif
#ifndef WX_RECPIPE_BS
#define WX_RECPIPE_BS 1 // back-substitution: same register pipeline
#endif
#ifndef WX_SPARSECLR
#define WX_SPARSECLR 1  // Z tile cleared once per iteration; per batch only rows of cameras that left the slot are zeroed
#endif
#ifndef WX_SCHOIST
#define WX_SCHOIST 0    // back-substitution: scratch row loaded at the top of the batch
#endif

// A 16-byte record is loaded with one LDG.128; its `orig` word is not used by the solve, and ptxas recycles a dead
// destination register at once -- the next write to it (a loop counter of the DMMA loop) then waits for the load
// (write-after-write on the long scoreboard).  A real, never-taken use where the record is consumed keeps the register
// allocated while the load is in flight: `orig` is an observation index, < 2^31 - 1 by the ABI's size limit.
__device__ __forceinline__ void keep_alive(const ObsRec& r) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.s32 p, %0, 0x7fffffff;\n\t@p trap;\n\t}" ::"r"(r.orig));
}

__device__ __forceinline__ int packed21(int r, int s) { return r * 6 - r * (r - 1) / 2 + (s - r); }

// ---- compile-time loop: f(integral_constant<int, I>) for I in [0, N)
template <int I, int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for<I + 1, N>(f);
  }
}

template <class Cfg, int GRP, int R>
constexpr bool row_used() {
  for (int q = 0; q < Cfg::TPW; ++q) {
    if (GRP * Cfg::TPW + q >= Cfg::T) continue;
    const int t = Cfg::tile_of(GRP, q);
    if (Tri<Cfg::NT>::mi(t) == R || Tri<Cfg::NT>::ni(t) == R) return true;
  }
  return false;
}

// acc += Zt-slice * Zt-slice^T for the tiles of warp group GRP over k-steps [kb, kb + KPER).
// zf = Zt + (lane & 3) * LDT + (lane >> 2): fragment (a and b alike) of tile row r at k-step k is
// zf[(4 k) * LDT + 8 r].  All offsets are immediates after unrolling.
template <class Cfg, int GRP>
__device__ __forceinline__ void syrk_group(const double* __restrict__ zf, int kb, double (&acc)[Cfg::TPW][2]) {
  constexpr int KPER = Cfg::KSTEPS / Cfg::KS;
  const double* zk = zf + (size_t)kb * 4 * Cfg::LDT;
#pragma unroll 2
  for (int ks = 0; ks < KPER; ++ks) {
    double fr[Cfg::NT];
    static_for<0, Cfg::NT>([&](auto r) {
      constexpr int R = decltype(r)::value;
      if constexpr (row_used<Cfg, GRP, R>()) fr[R] = zk[8 * R];
    });
    static_for<0, Cfg::TPW>([&](auto q) {
      constexpr int Q = decltype(q)::value;
      if constexpr (GRP * Cfg::TPW + Q < Cfg::T) {
        constexpr int t = Cfg::tile_of(GRP, Q);
        dmma884(acc[Q][0], acc[Q][1], fr[Tri<Cfg::NT>::mi(t)], fr[Tri<Cfg::NT>::ni(t)]);
      }
    });
    zk += 4 * Cfg::LDT;
  }
}

template <class Cfg, int GRP = 0>
__device__ __forceinline__ void syrk_dispatch(int grp, const double* __restrict__ zf, int kb,
                                              double (&acc)[Cfg::TPW][2]) {
  if constexpr (GRP < Cfg::G) {
    if (grp == GRP) syrk_group<Cfg, GRP>(zf, kb, acc);
    else syrk_dispatch<Cfg, GRP + 1>(grp, zf, kb, acc);
  }
}

template <class Cfg, int LOSS>
__global__ void __launch_bounds__(Cfg::NTHREADS, Cfg::MINB)
window_lm_kernel(const WinData d, const Params P) {
  constexpr int NTHREADS = Cfg::NTHREADS, NWARPS = Cfg::NWARPS, G = Cfg::G, KS = Cfg::KS, NT = Cfg::NT;
  constexpr int TPW = Cfg::TPW, TPP = Cfg::TPP, PB = Cfg::PB, LDT = Cfg::LDT, LDS = Cfg::LDS, MAXCAM = Cfg::MAXCAM;
  constexpr int PS = Cfg::PS;               // pose stride in shared memory (doubles)
  constexpr int CS = Cfg::CS, BS = Cfg::CS * Cfg::PB;   // CTAs per window; landmark stride between this CTA's batches
  extern __shared__ __align__(16) double smem[];
  double* pose_cur = smem;
  double* pose_tri = pose_cur + MAXCAM * PS;
  double* Bc = pose_tri + MAXCAM * PS;      // MAXCAM x 21 (upper, packed)
  double* gc = Bc + MAXCAM * 21;            // MAXCAM x 6
  double* Dc = gc + MAXCAM * 6;
  double* dvec = Dc + MAXCAM * 6;           // delta_c (6*nfree)
  double* gred = dvec + 128;
  double* red = gred + 128;                 // NWARPS x 32 reduction scratch
  double* scal = red + Cfg::NRED;           // 16 control scalars
  double* xbuf = scal + 16;                 // cluster exchange buffer (CS > 1)
  double* Z = xbuf + Cfg::XBUF;             // Zt tile (transposed), later Z Z^T tiles + S
  int* fidx = reinterpret_cast<int*>(Z + Cfg::ZS);   // free index or -1
  int* lo = fidx + MAXCAM;                  // # free cameras with index < c  (MAXCAM+1)
  int* iflag = lo + MAXCAM + 1;             // [0] step_bad  [1] accept  [2] stop

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int win = blockIdx.x / CS;
  int crank = 0;                            // rank of this CTA in the window's cluster
  if constexpr (CS > 1) crank = (int)cg::this_cluster().block_rank();
  const int c0 = d.cam_off[win], nc = d.cam_off[win + 1] - c0;
  const int p0 = d.pt_off[win], npw = d.pt_off[win + 1] - p0;

  for (int i = tid; i < nc * 12; i += NTHREADS) pose_cur[(i / 12) * PS + i % 12] = d.poses[(size_t)c0 * 12 + i];
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
  // first / last record of the window: unconditional record loads are clamped into the window's own stream (a record of
  // this window names one of its cameras, so whatever an idle lane computes from it stays finite and is weighted 0)
  const int rec_lo = d.pt_ptr[p0], rec_hi = max(d.pt_ptr[p0 + npw] - 1, rec_lo);
  const int n = 6 * nfree;              // reduced system size; row n of Z is the y row
  const int grp = warp % G, ksp = warp / G;

  double mu = P.init_mu, nu = 2.0, cost = 0.0, initial_cost = 0.0;
  int cur = 0, it = 0, n_acc = 0, n_rej = 0, status = B2BA_MAX_ITERS, ninv_cur = 0;
  double* const scratch = d.pt_scratch + (size_t)p0 * kPtScratch;
  const int slot = tid / TPP, sub = tid % TPP;
  const int np_round = ((npw + PB - 1) / PB) * PB;
  WinDbg dbgL{nc, npw, n};

  // ------------------------------------------------------------------ camera pass (a1+a2)
  // Linearisation of the camera side at the state (pose_sm, X): B_c (21) and g_c (6) of every camera into Bd / Bd +
  // MAXCAM * 21, cost and invalid count into scal[0] / scal[1] (cluster: summed over the CTAs in rank order).  It runs
  // once before the loop and then as the TRIAL EVALUATION of every step: the trial cost of a step is the cost of the
  // next linearisation point, so an accepted step (the common case) already holds its B_c, g_c and no separate
  // point-major trial pass re-projects the observations.
  auto camera_pass = [&](const double* pose_sm, const double* X, double* Bd) {
    double* gd = Bd + MAXCAM * 21;
      // one observation of camera pose q into the 21 + 6 + 2 sums (structural zeros of J_c skipped)
      auto cam_accum = [&](const ObsRec& rec, double x0, double x1, double x2, const Pose12& q, bool isfree, double* acc) {
        Geo o;
        geo_project<LOSS>(P, q, x0, x1, x2, rec.u, rec.v, o);
        if (!o.valid) { acc[28] += 1.0; return; }
        acc[27] = fma(0.5, o.rho, acc[27]);
        if (isfree) {
          double j0[6], j1[6];
          geo_jc(o, j0, j1);
          int k = 0;
#pragma unroll
          for (int r = 0; r < 6; ++r) {
            const double w0 = o.w * j0[r], w1 = o.w * j1[r];
#pragma unroll
            for (int s = r; s < 6; ++s, ++k) {
              if (r != 4 && s != 4) acc[k] = fma(w0, j0[s], acc[k]);
              if (r != 3 && s != 3) acc[k] = fma(w1, j1[s], acc[k]);
            }
            if (r != 4) acc[21 + r] = fma(w0, o.r0, acc[21 + r]);
            if (r != 3) acc[21 + r] = fma(w1, o.r1, acc[21 + r]);
          }
        }
      };
      // transposing butterfly: 31 shuffles instead of 29 x 5; lane l ends with the warp total of acc[l]
      auto warp_totals = [&](const double* acc) {
        double v[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) v[k] = k < 29 ? acc[k] : 0.0;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          const bool up = (lane & off) != 0;
#pragma unroll
          for (int j = 0; j < off; ++j) {
            const double keep = up ? v[j + off] : v[j];
            const double send = up ? v[j] : v[j + off];
            v[j] = keep + __shfl_xor_sync(0xffffffffu, send, off);
          }
        }
        return v[0];
      };
      double cst_tot = 0.0, ninv_tot = 0.0;   // live in thread 27 / 28
      if constexpr (CS > 1) {
        // latency path: the CTA holds only 1/CS of every camera's observations (a few per lane), so each WARP takes
        // whole cameras -- no cross-warp sum and no block barrier per camera
        double cst_w = 0.0, ninv_w = 0.0;     // lanes 27 / 28
        for (int c = warp; c < nc; c += NWARPS) {
          int b = d.cam_ptr[c0 + c], e = d.cam_ptr[c0 + c + 1];
          const int len = e - b;
          e = b + (int)(((long long)len * (crank + 1)) / CS);
          b = b + (int)(((long long)len * crank) / CS);
          Pose12 q;
          load_pose(pose_sm + PS * c, q);
          const bool isfree = fidx[c] >= 0;
          double acc[29];
#pragma unroll
          for (int k = 0; k < 29; ++k) acc[k] = 0.0;
          const ObsRec zrec{0.f, 0.f, 0, 0};
          ObsRec rec_n = zrec;
          if (b + lane < e) rec_n = d.cm[b + lane];
          for (int a = b + lane; a < e; a += 32) {
            const ObsRec rec = rec_n;
            if (a + 32 < e) { rec_n = d.cm[a + 32]; prefetch_l1(X + 3 * rec_n.idx); prefetch_l1(X + 3 * rec_n.idx + 2); }
            const double* Xp = X + 3 * rec.idx;
            cam_accum(rec, Xp[0], Xp[1], Xp[2], q, isfree, acc);
          }
          const double v = warp_totals(acc);
          if (lane < 21) Bd[c * 21 + lane] = v;
          else if (lane < 27) gd[c * 6 + lane - 21] = v;
          else if (lane == 27) cst_w += v;
          else if (lane == 28) ninv_w += v;
        }
        if (lane == 27) red[warp] = cst_w;
        if (lane == 28) red[32 + warp] = ninv_w;
        __syncthreads();
        if (tid == 27) for (int w = 0; w < NWARPS; ++w) cst_tot += red[w];
        if (tid == 28) for (int w = 0; w < NWARPS; ++w) ninv_tot += red[32 + w];
      } else {
      for (int c = 0; c < nc; ++c) {
        const int b = d.cam_ptr[c0 + c], e = d.cam_ptr[c0 + c + 1];
        Pose12 q;
        load_pose(pose_sm + PS * c, q);
        const bool isfree = fidx[c] >= 0;
        double acc[29];
#pragma unroll
        for (int k = 0; k < 29; ++k) acc[k] = 0.0;
        // software pipeline: records three iterations ahead in registers (unconditional loads, clamped into the
        // camera's run); the landmark a record names comes from L2 (the L1 left beside 4 x 55 KB of shared memory is
        // small), so it is L1-prefetched two iterations ahead and loaded into registers one iteration ahead
        if (b < e) {
          const int last = e - 1;
          ObsRec rec_n = d.cm[min(b + tid, last)], rec_nn = d.cm[min(b + tid + NTHREADS, last)];
          ObsRec rec_n3 = d.cm[min(b + tid + 2 * NTHREADS, last)];
          double xn0 = X[3 * rec_n.idx], xn1 = X[3 * rec_n.idx + 1], xn2 = X[3 * rec_n.idx + 2];
          prefetch_l1(X + 3 * rec_nn.idx); prefetch_l1(X + 3 * rec_nn.idx + 2);
          for (int a = b + tid; a < e; a += NTHREADS) {
            const ObsRec rec = rec_n;
            const double x0 = xn0, x1 = xn1, x2 = xn2;
            keep_alive(rec);
            rec_n = rec_nn;
            rec_nn = rec_n3;
            rec_n3 = d.cm[min(a + 3 * NTHREADS, last)];
            xn0 = X[3 * rec_n.idx]; xn1 = X[3 * rec_n.idx + 1]; xn2 = X[3 * rec_n.idx + 2];
            prefetch_l1(X + 3 * rec_nn.idx); prefetch_l1(X + 3 * rec_nn.idx + 2);
            cam_accum(rec, x0, x1, x2, q, isfree, acc);
          }
        }
        red[warp * 32 + lane] = warp_totals(acc);
        __syncthreads();
        if (tid < 29) {
          double v = 0.0;
          for (int w = 0; w < NWARPS; ++w) v += red[w * 32 + tid];
          if (tid < 21) Bd[c * 21 + tid] = v;
          else if (tid < 27) gd[c * 6 + tid - 21] = v;
          else if (tid == 27) cst_tot += v;
          else ninv_tot += v;
        }
        __syncthreads();
      }
      }
      if (tid == 27) scal[0] = cst_tot;
      if (tid == 28) scal[1] = ninv_tot;
      if constexpr (CS > 1) {
        // all-reduce {B_c, g_c, cost, n_invalid} over the cluster: partials go to xbuf, every CTA sums the
        // CS remote copies in rank order (deterministic and identical on all CTAs)
        __syncthreads();
        constexpr int NB = MAXCAM * 27;          // B and g are contiguous (Bc | gc, and their trial copies alike)
        for (int i = tid; i < NB; i += NTHREADS) xbuf[i] = Bd[i];
        if (tid < 2) xbuf[NB + tid] = scal[tid];
        cg::cluster_group cl = cg::this_cluster();
        cl.sync();
        for (int i = tid; i < NB + 2; i += NTHREADS) {
          double v = 0.0;
#pragma unroll
          for (int r = 0; r < CS; ++r) v += cl.map_shared_rank(xbuf, r)[i];
          if (i < NB) Bd[i] = v; else scal[i - NB] = v;
        }
        // no second barrier: the next writes into xbuf[0, NB + 2) are the reduce-scattered tiles of the next point
        // pass, and those come after that pass's first cluster barrier, which every CTA reaches after these reads
      }
    __syncthreads();
  };
  double* const Bt = Z;                       // trial B_c | g_c (the tile memory is free between the solve and the next point pass)
  int fresh_lin = 1;                          // the linearisation at hand has not been solved with yet
  camera_pass(pose_cur, d.pts[0] + (size_t)p0 * 3, Bc);
  for (int i = tid; i < nc * 6; i += NTHREADS) Dc[i] = clampd(Bc[(i / 6) * 21 + packed21(i % 6, i % 6)]);
  cost = scal[0];
  ninv_cur = (int)scal[1];
  initial_cost = cost;
  __syncthreads();

  for (; it < P.max_iters; ++it) {
    const double* X = d.pts[cur] + (size_t)p0 * 3;
    double* Xt = d.pts[cur ^ 1] + (size_t)p0 * 3;
    const bool dbg = (d.dbg != nullptr) && it == 0;

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
    double gmax = 0.0;
    const double imu = 1.0 / mu;
    const double* zf = Z + (lane & 3) * LDT + (lane >> 2);   // DMMA fragment base of this lane
    double* const zcol = Z + (size_t)3 * slot * LDT;          // this landmark's 3 columns: 3*LDT contiguous doubles
    const ObsRec zrec{0.f, 0.f, 0, 0};
    unsigned prev_rows = 0;                   // free cameras (+ bit nfree: the y row) this slot wrote in the previous batch
    if constexpr (WX_SPARSECLR) {
      // the tile memory held S and the Cholesky factor: clear it once; afterwards a slot only zeroes the rows of cameras
      // that saw its previous landmark but do not see the present one
      double2* z2 = reinterpret_cast<double2*>(Z);
      const double2 zz = make_double2(0.0, 0.0);
      for (int i = tid; i < Cfg::NCOL * LDT / 2; i += NTHREADS) z2[i] = zz;
      __syncthreads();
    }
    int a0n = 0, a1n = 0, a0nn = 0, a1nn = 0;
    double Xn0 = 0, Xn1 = 0, Xn2 = 0;
    ObsRec rn0 = zrec, rn1 = zrec;            // this lane's first two records of the NEXT batch (registers)
    const int bfirst = crank * PB;            // this CTA takes batches bfirst, bfirst + BS, ...
    // landmark slots are in schedule order (decreasing observation count, set up by the index builder), so the
    // landmarks sharing a warp do the same number of rounds and all per-landmark arrays are contiguous
    if (bfirst + slot < npw) {
      const int pf = bfirst + slot;
      a0n = d.pt_ptr[p0 + pf]; a1n = d.pt_ptr[p0 + pf + 1];
      Xn0 = X[3 * pf]; Xn1 = X[3 * pf + 1]; Xn2 = X[3 * pf + 2];
      rn0 = d.pm[min(max(a0n, rec_lo) + sub, rec_hi)];
      rn1 = d.pm[min(max(a0n, rec_lo) + sub + TPP, rec_hi)];
    }
    if (bfirst + BS + slot < npw) { a0nn = d.pt_ptr[p0 + bfirst + BS + slot]; a1nn = d.pt_ptr[p0 + bfirst + BS + slot + 1]; }
    int a0n3 = 0, a1n3 = 0;
    if (bfirst + 2 * BS + slot < npw) { a0n3 = d.pt_ptr[p0 + bfirst + 2 * BS + slot]; a1n3 = d.pt_ptr[p0 + bfirst + 2 * BS + slot + 1]; }
    // software pipeline of the record stream: while batch b runs its DMMA, the records of batch b + 1 are loaded into
    // registers (L1 hits: their lines were prefetched during batch b - 1), the lines of batch b + 2 are L1-prefetched
    // by the landmark's own lanes (a landmark's records are one contiguous run of <= 3 lines), and the landmark
    // offsets of batch b + 3 are loaded -- every address is in a register one batch before it is needed
    auto prefetch_records = [&](int s0_, int s1_) {
      const int a = (s0_ & ~7) + 8 * sub;
      if (a < s1_) prefetch_l1(d.pm + a);
    };
    prefetch_records(a0nn, a1nn);

    for (int b0 = bfirst; b0 < npw; b0 += BS) {
      const int p = b0 + slot;
      const bool act = p < npw;
      const int a0 = a0n, a1 = a1n;
      const double X0 = Xn0, X1 = Xn1, X2 = Xn2;
      ObsRec rec0 = rn0, rec1 = rn1;
      keep_alive(rec0); keep_alive(rec1);
      if constexpr (!WX_RECPIPE_PT) {
        rec0 = zrec; rec1 = zrec;
        if (a0 + sub < a1) rec0 = d.pm[a0 + sub];
        if (a0 + sub + TPP < a1) rec1 = d.pm[a0 + sub + TPP];
      }
      // clear the column block (rows of cameras that do not see the landmark must be zero; the
      // block is contiguous, so this is TPP lanes x STS.128)
      if constexpr (!WX_SPARSECLR || WX_STASH) {
        double2* zc2 = reinterpret_cast<double2*>(zcol);
        const double2 zz = make_double2(0.0, 0.0);
#pragma unroll 4
        for (int i = sub; i < 3 * LDT / 2; i += TPP) zc2[i] = zz;
        __syncwarp();   // column block cleared by all TPP lanes before anyone writes into it
      }
      // loop 1: C_p, g_p.  A lane takes observations a0+sub, a0+sub+TPP, ...; they are handled two at a
      // time in one basic block (independent dependency chains interleave); `rounds` is warp-uniform.  What loop 2
      // needs of a free camera's observation -- w J_p (2x3) and the 10 non-zero entries of J_c -- is parked in the 18
      // doubles of the column block that will hold this observation's Z rows, so loop 2 projects nothing.
      const int rounds = (int)((__reduce_max_sync(0xffffffffu, (unsigned)(a1 - a0)) + TPP - 1) / TPP);
      double C[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
      auto accum = [&](const ObsRec& rec, bool lv) -> int {
        const int f = lv ? fidx[rec.idx] : -1;
        Pose12 q;
        load_pose(pose_cur + PS * rec.idx, q);
        Geo o;
        geo_project<LOSS>(P, q, X0, X1, X2, rec.u, rec.v, o);
        const double w = lv ? o.w : 0.0;
        double jp[6];
        geo_jp(q, o, jp);
        const double u0 = w * jp[0], u1 = w * jp[1], u2 = w * jp[2];
        const double v0 = w * jp[3], v1 = w * jp[4], v2 = w * jp[5];
        C[0] = fma(u0, jp[0], fma(v0, jp[3], C[0])); C[1] = fma(u0, jp[1], fma(v0, jp[4], C[1]));
        C[2] = fma(u0, jp[2], fma(v0, jp[5], C[2])); C[3] = fma(u1, jp[1], fma(v1, jp[4], C[3]));
        C[4] = fma(u1, jp[2], fma(v1, jp[5], C[4])); C[5] = fma(u2, jp[2], fma(v2, jp[5], C[5]));
        g[0] = fma(u0, o.r0, fma(v0, o.r1, g[0]));
        g[1] = fma(u1, o.r0, fma(v1, o.r1, g[1]));
        g[2] = fma(u2, o.r0, fma(v2, o.r1, g[2]));
        if constexpr (WX_STASH) {
        double j0[6], j1[6];
        geo_jc(o, j0, j1);
        if (f >= 0) {
          double* zr = zcol + 6 * f;
          double2* c0 = reinterpret_cast<double2*>(zr);
          double2* c1 = reinterpret_cast<double2*>(zr + LDT);
          double2* c2 = reinterpret_cast<double2*>(zr + 2 * LDT);
          c0[0] = make_double2(u0, u1); c0[1] = make_double2(u2, v0); c0[2] = make_double2(v1, v2);
          c1[0] = make_double2(j0[0], j0[1]); c1[1] = make_double2(j0[2], j0[3]); c1[2] = make_double2(j0[5], j1[0]);
          c2[0] = make_double2(j1[1], j1[2]); c2[1] = make_double2(j1[4], j1[5]);
        }
        }
        return f;
      };
      int f0 = -1, f1 = -1;
      if (rounds > 1) {
        f0 = accum(rec0, a0 + sub < a1);
        f1 = accum(rec1, a0 + sub + TPP < a1);
      } else if (rounds == 1) {
        f0 = accum(rec0, a0 + sub < a1);
      }
      unsigned rows = (f0 >= 0 ? 1u << f0 : 0u) | (f1 >= 0 ? 1u << f1 : 0u);
      for (int r = 2; r < rounds; r += 2) {
        const int aA = a0 + sub + r * TPP, aB = aA + TPP;
        const bool lvA = aA < a1, lvB = aB < a1;
        ObsRec recA = zrec, recB = zrec;
        if (lvA) recA = d.pm[aA];
        if (r + 1 < rounds) {
          if (lvB) recB = d.pm[aB];
          const int fa = accum(recA, lvA);
          const int fb = accum(recB, lvB);
          rows |= (fa >= 0 ? 1u << fa : 0u) | (fb >= 0 ? 1u << fb : 0u);
        } else {
          const int fa = accum(recA, lvA);
          rows |= (fa >= 0 ? 1u << fa : 0u);
        }
      }
      if constexpr (WX_SPARSECLR && !WX_STASH) {
        if (act) rows |= 1u << nfree;           // the y row (rows n ...)
#pragma unroll
        for (int m = TPP >> 1; m > 0; m >>= 1) rows |= __shfl_xor_sync(0xffffffffu, rows, m);
        unsigned cl = prev_rows & ~rows;
        prev_rows = rows;
        const double2 zz = make_double2(0.0, 0.0);
        for (int k = 0; cl; ++k) {
          const int f = __ffs(cl) - 1;
          cl &= cl - 1;
          if ((k & (TPP - 1)) == sub) {
            double* zr = zcol + 6 * f;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
              double2* dst = reinterpret_cast<double2*>(zr + j * LDT);
              dst[0] = zz; dst[1] = zz; dst[2] = zz;
            }
          }
        }
        __syncwarp();   // rows zeroed before another lane of the landmark writes them
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
        const double Cd[6] = {fma(D0, imu, C[0]), C[1], C[2], fma(D1, imu, C[3]), C[4], fma(D2, imu, C[5])};
        if (!sym3_inv_factor_fast(Cd, F)) {
#pragma unroll
          for (int k = 0; k < 6; ++k) F[k] = 0.0;
          if (act) iflag[0] = 1;
        }
      }
      gmax = fmax(gmax, fmax(fabs(g[0]), fmax(fabs(g[1]), fabs(g[2]))));
      if (sub == 0 && act) {
        zcol[n] = F[0] * g[0];
        zcol[LDT + n] = fma(F[1], g[0], F[3] * g[1]);
        zcol[2 * LDT + n] = fma(F[2], g[0], fma(F[4], g[1], F[5] * g[2]));
        double2* sc = reinterpret_cast<double2*>(scratch + (size_t)p * kPtScratch);
        sc[0] = make_double2(F[0], F[1]); sc[1] = make_double2(F[2], F[3]); sc[2] = make_double2(F[4], F[5]);
        sc[3] = make_double2(g[0], g[1]); sc[4] = make_double2(g[2], D0); sc[5] = make_double2(D1, D2);
        if (dbg) {
#pragma unroll
          for (int k = 0; k < 6; ++k) d.dbg[dbgL.off_C() + (size_t)p * 6 + k] = C[k];
#pragma unroll
          for (int k = 0; k < 3; ++k) d.dbg[dbgL.off_gp() + (size_t)p * 3 + k] = g[k];
        }
      }
      // loop 2: Z rows of free camera f = E_k F = J_c^T ((w J_p) F), 6 x 3, from the parked values, in place
      auto emit = [&](int f) {
        const bool st = f >= 0;
        double* zr = zcol + 6 * (st ? f : 0);
        double2* c0 = reinterpret_cast<double2*>(zr);
        double2* c1 = reinterpret_cast<double2*>(zr + LDT);
        double2* c2 = reinterpret_cast<double2*>(zr + 2 * LDT);
        const double2 wa = c0[0], wb = c0[1], wc = c0[2];
        const double2 ja = c1[0], jb = c1[1], jc_ = c1[2];
        const double2 jd = c2[0], je = c2[1];
        const double w0 = wa.x, w1 = wa.y, w2 = wb.x, w3 = wb.y, w4 = wc.x, w5 = wc.y;
        // G = (w J_p) F  (2 x 3; F upper triangular)
        const double G0[3] = {w0 * F[0], fma(w0, F[1], w1 * F[3]), fma(w0, F[2], fma(w1, F[4], w2 * F[5]))};
        const double G1[3] = {w3 * F[0], fma(w3, F[1], w4 * F[3]), fma(w3, F[2], fma(w4, F[4], w5 * F[5]))};
        // J_c rows: j0 = (ja.x ja.y jb.x jb.y 0 jc.x), j1 = (jc.y jd.x jd.y 0 je.x je.y)
        if (st) {
          double2* dst[3] = {c0, c1, c2};
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            dst[j][0] = make_double2(fma(ja.x, G0[j], jc_.y * G1[j]), fma(ja.y, G0[j], jd.x * G1[j]));
            dst[j][1] = make_double2(fma(jb.x, G0[j], jd.y * G1[j]), jb.y * G0[j]);
            dst[j][2] = make_double2(je.x * G1[j], fma(jc_.x, G0[j], je.y * G1[j]));
          }
        }
      };
      // (WX_STASH == 0: re-project instead)
      auto emit_rec = [&](const ObsRec& rec, int f) {
        const bool st = f >= 0;
        Pose12 q;
        load_pose(pose_cur + PS * rec.idx, q);
        Geo o;
        geo_project<LOSS>(P, q, X0, X1, X2, rec.u, rec.v, o);
        double jp[6];
        geo_jp(q, o, jp);
        const double w0 = o.w * jp[0], w1 = o.w * jp[1], w2 = o.w * jp[2];
        const double w3 = o.w * jp[3], w4 = o.w * jp[4], w5 = o.w * jp[5];
        const double G0[3] = {w0 * F[0], fma(w0, F[1], w1 * F[3]), fma(w0, F[2], fma(w1, F[4], w2 * F[5]))};
        const double G1[3] = {w3 * F[0], fma(w3, F[1], w4 * F[3]), fma(w3, F[2], fma(w4, F[4], w5 * F[5]))};
        double j0[6], j1[6];
        geo_jc(o, j0, j1);
        if (st) {
          double* zr = zcol + 6 * f;
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            double2* dst = reinterpret_cast<double2*>(zr + j * LDT);
            dst[0] = make_double2(fma(j0[0], G0[j], j1[0] * G1[j]), fma(j0[1], G0[j], j1[1] * G1[j]));
            dst[1] = make_double2(fma(j0[2], G0[j], j1[2] * G1[j]), j0[3] * G0[j]);
            dst[2] = make_double2(j1[4] * G1[j], fma(j0[5], G0[j], j1[5] * G1[j]));
          }
        }
      };
      if constexpr (WX_STASH) {
        if (rounds > 1) {
          emit(f0);
          emit(f1);
        } else if (rounds == 1) {
          emit(f0);
        }
      } else {
        if (rounds > 1) {
          emit_rec(rec0, f0);
          emit_rec(rec1, f1);
        } else if (rounds == 1) {
          emit_rec(rec0, f0);
        }
      }
      for (int r = 2; r < rounds; ++r) {
        const int a = a0 + sub + r * TPP;
        int f = -1;
        ObsRec rec = zrec;
        if (a < a1) { rec = d.pm[a]; f = fidx[rec.idx]; }
        if constexpr (WX_STASH) emit(f); else emit_rec(rec, f);
      }
      __syncthreads();
      // the next batch's landmark and first two records (registers) and the lines of the batch after it (L1
      // prefetch: a landmark's records are one contiguous run of <= 2-3 lines) are fetched while the tensor pipe works
      {
        const int pn = p + BS;
        a0n = a0nn; a1n = a1nn;
        if constexpr (WX_RECPIPE_PT) {
          // unconditional loads from clamped addresses (no zero-fill of the destination, which would have to wait for
          // the previous load into the same registers); validity is decided by the offsets when the record is used
          rn0 = d.pm[min(max(a0n, rec_lo) + sub, rec_hi)];
          rn1 = d.pm[min(max(a0n, rec_lo) + sub + TPP, rec_hi)];
        }
        if (pn < npw) { Xn0 = X[3 * pn]; Xn1 = X[3 * pn + 1]; Xn2 = X[3 * pn + 2]; }
        a0nn = a0n3; a1nn = a1n3;
        prefetch_records(a0nn, a1nn);
        a0n3 = 0; a1n3 = 0;
        if (pn + 2 * BS < npw) { a0n3 = d.pt_ptr[p0 + pn + 2 * BS]; a1n3 = d.pt_ptr[p0 + pn + 2 * BS + 1]; }
      }
      // Z Z^T accumulate: DMMA over this batch's 3*PB columns (this warp's k-slice)
      syrk_dispatch<Cfg>(grp, zf, ksp * (Cfg::KSTEPS / KS), acc);
      __syncthreads();
    }

    // ------------------------------------------------------------------ S = B' - Z Z^T  (a4)
    // the Z tile is dead now: its memory takes this CTA's Z Z^T tiles (SP, T x 64, 8x8 row-major each) and
    // the assembled system S (n x n, leading dim LDS)
    double* SP = Z;
    double* S = Z + Cfg::SPART;
#pragma unroll
    for (int s = 0; s < KS; ++s) {
      if (ksp == s) {
        static_for<0, G>([&](auto gg) {
          constexpr int GG = decltype(gg)::value;
          if (grp == GG) {
            static_for<0, TPW>([&](auto qq) {
              constexpr int Q = decltype(qq)::value;
              if constexpr (GG * TPW + Q < Cfg::T) {
                constexpr int t = Cfg::tile_of(GG, Q);
                double* dst = SP + t * 64 + (lane >> 2) * 8 + 2 * (lane & 3);
                if (s == 0) { dst[0] = acc[Q][0]; dst[1] = acc[Q][1]; }
                else { dst[0] += acc[Q][0]; dst[1] += acc[Q][1]; }
              }
            });
          }
        });
      }
      __syncthreads();
    }
    if constexpr (CS > 1) {
      // a landmark block that failed to factor in ANY CTA invalidates the step for the whole window
      if (tid == 0) xbuf[Cfg::XDATA + 8] = iflag[0] ? 1.0 : 0.0;
      cg::cluster_group cl = cg::this_cluster();
      cl.sync();                                 // all CTAs' tiles (and flags) are in place
      if (tid == 0) {
        bool bad = false;
        for (int r = 0; r < CS; ++r) bad = bad || cl.map_shared_rank(xbuf, r)[Cfg::XDATA + 8] != 0.0;
        if (bad) iflag[0] = 1;
      }
      // reduce-scatter: this CTA sums its XCHUNK-entry share of the tiles over all CTAs (rank order) ...
      for (int e = tid; e < Cfg::XCHUNK; e += NTHREADS) {
        const int off = crank * Cfg::XCHUNK + e;
        double v = 0.0;
        if (off < Cfg::SPART) {
#pragma unroll
          for (int r = 0; r < CS; ++r) v += cl.map_shared_rank(SP, r)[off];
        }
        xbuf[e] = v;
      }
      cl.sync();                                 // ... and every CTA then gathers the sums from their owners:
      // independent remote loads, a few per thread, into its own tile memory (no peer reads that any more), so the
      // assembly below works on local shared memory in every variant
      for (int e = tid; e < Cfg::SPART; e += NTHREADS) SP[e] = cl.map_shared_rank(xbuf, e / Cfg::XCHUNK)[e % Cfg::XCHUNK];
      __syncthreads();
    }
    // (Z Z^T)[i][j] for i <= j: one tile entry (cluster: summed over the CTAs in rank order)
    auto zzt = [&](int i, int j) -> double {
      const int mi = i >> 3, ni = j >> 3;
      const int off = (mi * NT - mi * (mi - 1) / 2 + (ni - mi)) * 64 + (i & 7) * 8 + (j & 7);
      return SP[off];
    };
    for (int idx = tid; idx < n * n; idx += NTHREADS) {
      const int i = idx / n, j = idx - i * n;
      if (i > j) continue;
      double v = -zzt(i, j);
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
      gred[tid] = gc[c * 6 + tid - 6 * fi] - zzt(tid, n);
    }
    __syncthreads();
    if (dbg) {
      for (int idx = tid; idx < n * n; idx += NTHREADS) d.dbg[dbgL.off_S() + idx] = S[(idx / n) * LDS + idx % n];
      for (int i = tid; i < n; i += NTHREADS) d.dbg[dbgL.off_gred() + i] = gred[i];
      __syncthreads();
    }

    // ------------------------------------------------------------------ Cholesky + solve (a5)
    // right-looking Cholesky in 6-column (camera) blocks, lower triangle of S, 2 barriers per block.  The right-hand side
    // -g_red rides along as row n of S: panel solves and trailing updates treat it like any other row, so the forward
    // substitution y = L^-1 b falls out of the factorisation in parallel and only the backward solve stays serial
    for (int k = tid; k < n; k += NTHREADS) S[n * LDS + k] = -gred[k];
    __syncthreads();
    for (int j0 = 0; j0 < n; j0 += 6) {
      // every thread factors the 6x6 diagonal block redundantly in registers: L (packed lower, row-major)
      double L[6][6], inv[6];
      bool bad = false;
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c)
          if (c <= r) L[r][c] = S[(j0 + r) * LDS + j0 + c];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        double dj = L[c][c];
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (k < c) dj = fma(-L[c][k], L[c][k], dj);
        if (!(dj > 0.0) || !isfinite(dj)) { bad = true; dj = 1.0; }
        inv[c] = rsqrt_pos(dj);
        L[c][c] = dj * inv[c];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          if (r > c) {
            double v = L[r][c];
#pragma unroll
            for (int k = 0; k < 6; ++k)
              if (k < c) v = fma(-L[r][k], L[c][k], v);
            L[r][c] = v * inv[c];
          }
        }
      }
      if (bad && tid == 0) iflag[0] = 1;
      const int m = n - j0 - 6;
      // panel: row i of L21 = A21 row * L11^-T, one thread per row (row m = the right-hand side)
      for (int t = tid; t <= m; t += NTHREADS) {
        double* row = S + (j0 + 6 + t) * LDS + j0;
        double x[6];
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          double v = row[c];
#pragma unroll
          for (int k = 0; k < 6; ++k)
            if (k < c) v = fma(-x[k], L[c][k], v);
          x[c] = v * inv[c];
        }
#pragma unroll
        for (int c = 0; c < 6; ++c) row[c] = x[c];
      }
      __syncthreads();
      if (tid == 0) {
#pragma unroll
        for (int r = 0; r < 6; ++r) {
#pragma unroll
          for (int c = 0; c < 6; ++c)
            if (c <= r) S[(j0 + r) * LDS + j0 + c] = L[r][c];
          red[j0 + r] = inv[r];               // 1 / L_rr for the triangular solves (red is free here)
        }
      }
      // trailing update A22 -= L21 L21^T (lower triangle), work items = (row, 6-column block) pairs
      const int npairs = (m / 6) * (m + 6) / 2 + m / 6;   // sum over column blocks cb of (m - cb + 1): rows cb .. m
      for (int id = tid; id < npairs; id += NTHREADS) {
        int cb = 0, rem = id, len = m + 1;
        while (rem >= len) { rem -= len; len -= 6; cb += 6; }
        const int ii = cb + rem;
        double* ri = S + (j0 + 6 + ii) * LDS + j0;
        double li[6];
#pragma unroll
        for (int c = 0; c < 6; ++c) li[c] = ri[c];
#pragma unroll
        for (int cc = 0; cc < 6; ++cc) {
          const int kc = cb + cc;
          if (kc <= ii) {
            const double* lk = S + (j0 + 6 + kc) * LDS + j0;
            double v = ri[6 + kc];
#pragma unroll
            for (int c = 0; c < 6; ++c) v = fma(-li[c], lk[c], v);
            ri[6 + kc] = v;
          }
        }
      }
      __syncthreads();
    }
    __syncthreads();
    if (warp == 0) {
      // L^T x = y (y = row n of S, see above) with the vector in registers (row k of lane k % 32), one shuffle per step.
      // Every lane keeps its unknowns pre-scaled by its own 1 / L_kk (x_k = d_k y_k - sum (d_k L_ik) x_i), so the
      // dependent chain of a step is shuffle + FMA only; the coefficients of step i - 1 are loaded before step i is consumed.
      constexpr int RPL = Cfg::NR / 32 + (Cfg::NR % 32 ? 1 : 0);
      double x[RPL], di[RPL], cn[RPL];
#pragma unroll
      for (int q = 0; q < RPL; ++q) {
        const int k = lane + 32 * q;
        di[q] = k < n ? red[k] : 0.0;
        x[q] = k < n ? S[n * LDS + k] * di[q] : 0.0;
        cn[q] = (k < n - 1) ? S[(n - 1) * LDS + k] * di[q] : 0.0;
      }
      for (int i = n - 1; i >= 0; --i) {
        double c[RPL];
#pragma unroll
        for (int q = 0; q < RPL; ++q) {
          const int k = lane + 32 * q;
          c[q] = cn[q];
          cn[q] = (i > 0 && k < i - 1) ? S[(i - 1) * LDS + k] * di[q] : 0.0;
        }
        double xi = 0.0;
#pragma unroll
        for (int q = 0; q < RPL; ++q)
          if ((i >> 5) == q) xi = __shfl_sync(0xffffffffu, x[q], i & 31);
#pragma unroll
        for (int q = 0; q < RPL; ++q) x[q] = fma(-c[q], xi, x[q]);
      }
#pragma unroll
      for (int q = 0; q < RPL; ++q) {
        const int k = lane + 32 * q;
        if (k < n) dvec[k] = x[q];
      }
    }
    __syncthreads();
    // trial cameras + camera part of the model decrease
    if (tid < nc) {
      const int f = fidx[tid];
      if (f >= 0) retract_pose(pose_cur + PS * tid, dvec + 6 * f, pose_tri + PS * tid);
      else
        for (int k = 0; k < 12; ++k) pose_tri[PS * tid + k] = pose_cur[PS * tid + k];
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
    double s_dDd = 0.0, s_gd = 0.0, s_dn2 = 0.0, s_xn2 = 0.0;
    if (step_ok) {
      const int qfirst = crank * PB + slot;
      const ObsRec zrec{0.f, 0.f, 0, 0};
      int a0n = 0, a1n = 0, a0nn = 0, a1nn = 0;
      ObsRec rn0 = zrec, rn1 = zrec;          // this lane's first two records of the NEXT batch (registers)
      if (qfirst < npw) {
        a0n = d.pt_ptr[p0 + qfirst]; a1n = d.pt_ptr[p0 + qfirst + 1];
        rn0 = d.pm[min(max(a0n, rec_lo) + sub, rec_hi)];
        rn1 = d.pm[min(max(a0n, rec_lo) + sub + TPP, rec_hi)];
      }
      if (qfirst + BS < npw) { a0nn = d.pt_ptr[p0 + qfirst + BS]; a1nn = d.pt_ptr[p0 + qfirst + BS + 1]; }
      int a0n3 = 0, a1n3 = 0;
      if (qfirst + 2 * BS < npw) { a0n3 = d.pt_ptr[p0 + qfirst + 2 * BS]; a1n3 = d.pt_ptr[p0 + qfirst + 2 * BS + 1]; }
      // same pipeline as the point pass: records one batch ahead in registers, their lines two batches ahead in L1
      // (prefetched by the landmark's own lanes), the landmark offsets three batches ahead
      auto prefetch_records = [&](int s0_, int s1_) {
        const int a = (s0_ & ~7) + 8 * sub;
        if (a < s1_) prefetch_l1(d.pm + a);
      };
      prefetch_records(a0nn, a1nn);
      for (int p = qfirst; p < np_round; p += BS) {
        const bool act = p < npw;
        const int a0 = a0n, a1 = a1n;
        ObsRec rec0 = rn0, rec1 = rn1;
        keep_alive(rec0); keep_alive(rec1);
        if constexpr (!WX_RECPIPE_BS) {
          rec0 = zrec; rec1 = zrec;
          if (a0 + sub < a1) rec0 = d.pm[a0 + sub];
          if (a0 + sub + TPP < a1) rec1 = d.pm[a0 + sub + TPP];
        }
        double X0 = 0, X1 = 0, X2 = 0;
        double2 s0 = make_double2(0, 0), s1 = s0, s2 = s0, s3 = s0, s4 = s0, s5 = s0;
        if (act) {
          X0 = X[3 * p]; X1 = X[3 * p + 1]; X2 = X[3 * p + 2];
          if constexpr (WX_SCHOIST) {
            const double2* sc = reinterpret_cast<const double2*>(scratch + (size_t)p * kPtScratch);
            s0 = sc[0]; s1 = sc[1]; s2 = sc[2]; s3 = sc[3]; s4 = sc[4]; s5 = sc[5];
          }
        }
        // next batch: offsets from registers, first two records into registers (L1 hits); offsets of the batch after it
        a0n = a0nn; a1n = a1nn;
        if constexpr (WX_RECPIPE_BS) {
          rn0 = d.pm[min(max(a0n, rec_lo) + sub, rec_hi)];
          rn1 = d.pm[min(max(a0n, rec_lo) + sub + TPP, rec_hi)];
        }
        a0nn = a0n3; a1nn = a1n3;
        prefetch_records(a0nn, a1nn);
        a0n3 = 0; a1n3 = 0;
        if (p + 3 * BS < npw) { a0n3 = d.pt_ptr[p0 + p + 3 * BS]; a1n3 = d.pt_ptr[p0 + p + 3 * BS + 1]; }
        {   // L1 prefetch: landmarks and scratch rows of the next batch
          const int bn = p - slot + BS, bnn = min(bn + PB, npw);
          if (bn < npw) {
            if (tid * 16 < 3 * (bnn - bn)) prefetch_l1(X + 3 * bn + tid * 16);
            if (tid * 16 < kPtScratch * (bnn - bn)) prefetch_l1(scratch + (size_t)bn * kPtScratch + tid * 16);
          }
        }
        const int rounds = (int)((__reduce_max_sync(0xffffffffu, (unsigned)(a1 - a0)) + TPP - 1) / TPP);
        double v[3] = {0, 0, 0};
        auto etd = [&](const ObsRec& rec, bool lv) {
          const int f = fidx[rec.idx];
          const bool use = lv && f >= 0;
          Pose12 q;
          load_pose(pose_cur + PS * rec.idx, q);
          Geo o;
          geo_project<LOSS>(P, q, X0, X1, X2, rec.u, rec.v, o);
          double jp[6], j0[6], j1[6];
          geo_jp(q, o, jp);
          geo_jc(o, j0, j1);
          // E^T dc = w J_p^T (J_c dc)
          const double* dc = dvec + 6 * (f < 0 ? 0 : f);
          const double w = use ? o.w : 0.0;
          const double t0 = w * fma(j0[0], dc[0], fma(j0[1], dc[1], fma(j0[2], dc[2], fma(j0[3], dc[3], j0[5] * dc[5]))));
          const double t1 = w * fma(j1[0], dc[0], fma(j1[1], dc[1], fma(j1[2], dc[2], fma(j1[4], dc[4], j1[5] * dc[5]))));
#pragma unroll
          for (int j = 0; j < 3; ++j) v[j] = fma(jp[j], t0, fma(jp[3 + j], t1, v[j]));
        };
        if (rounds > 1) {
          etd(rec0, a0 + sub < a1);
          etd(rec1, a0 + sub + TPP < a1);
        } else if (rounds == 1) {
          etd(rec0, a0 + sub < a1);
        }
        for (int r = 2; r < rounds; r += 2) {
          const int aA = a0 + sub + r * TPP, aB = aA + TPP;
          const bool lvA = aA < a1, lvB = aB < a1;
          ObsRec recA = zrec, recB = zrec;
          if (lvA) recA = d.pm[aA];
          if (r + 1 < rounds) {
            if (lvB) recB = d.pm[aB];
            etd(recA, lvA);
            etd(recB, lvB);
          } else {
            etd(recA, lvA);
          }
        }
#pragma unroll
        for (int m = TPP >> 1; m > 0; m >>= 1) {
#pragma unroll
          for (int k = 0; k < 3; ++k) v[k] += __shfl_xor_sync(0xffffffffu, v[k], m);
        }
        if (act) {
          if constexpr (!WX_SCHOIST) {
            const double2* sc = reinterpret_cast<const double2*>(scratch + (size_t)p * kPtScratch);
            s0 = sc[0]; s1 = sc[1]; s2 = sc[2]; s3 = sc[3]; s4 = sc[4]; s5 = sc[5];
          }
          const double F0 = s0.x, F1 = s0.y, F2 = s1.x, F3 = s1.y, F4 = s2.x, F5 = s2.y;
          const double g0 = s3.x, g1 = s3.y, g2 = s4.x;
          const double t0 = g0 + v[0], t1 = g1 + v[1], t2 = g2 + v[2];
          // delta_p = -F F^T t
          const double u0 = F0 * t0, u1 = fma(F1, t0, F3 * t1), u2 = fma(F2, t0, fma(F4, t1, F5 * t2));
          const double d0 = -fma(F0, u0, fma(F1, u1, F2 * u2)), d1 = -fma(F3, u1, F4 * u2), d2 = -(F5 * u2);
          if (sub == 0) {
            Xt[3 * p] = X0 + d0; Xt[3 * p + 1] = X1 + d1; Xt[3 * p + 2] = X2 + d2;
            s_dDd += d0 * d0 * s4.y + d1 * d1 * s5.x + d2 * d2 * s5.y;
            s_gd += d0 * g0 + d1 * g1 + d2 * g2;
            s_dn2 += d0 * d0 + d1 * d1 + d2 * d2;
            s_xn2 += X0 * X0 + X1 * X1 + X2 * X2;
            if (dbg) { double* dp = d.dbg + dbgL.off_dp() + (size_t)p * 3; dp[0] = d0; dp[1] = d1; dp[2] = d2; }
          }
        }
      }
    }
    {
      double vals[4] = {s_dDd, s_gd, s_dn2, s_xn2};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const double v = warp_sum(vals[k]);
        if (lane == 0) red[warp * 32 + k] = v;
      }
      double gm = gmax;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) gm = fmax(gm, __shfl_xor_sync(0xffffffffu, gm, o));
      if (lane == 0) red[warp * 32 + 4] = gm;
    }
    __syncthreads();                             // the sums; the trial landmarks are visible to the whole CTA
    // landmark-part sums of this CTA (fixed order over the warps) -> scal[9..13]; in a cluster also -> xbuf: the barrier
    // that makes every CTA's trial landmarks visible to the cluster publishes them, and every CTA adds the CS copies in
    // rank order after the trial evaluation
    if (tid == 0) {
      double t[5] = {0, 0, 0, 0, 0};
      for (int w = 0; w < NWARPS; ++w) {
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] += red[w * 32 + k];
        t[4] = fmax(t[4], red[w * 32 + 4]);
      }
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        scal[9 + k] = t[k];
        if constexpr (CS > 1) xbuf[Cfg::XDATA + k] = t[k];
      }
    }
    if constexpr (CS > 1) cg::this_cluster().sync();
    else __syncthreads();                        // `red` is the camera pass's scratch from here on

    // ------------------------------------------------------------------ trial evaluation = next linearisation (a1+a2, a7)
    if (step_ok) camera_pass(pose_tri, Xt, Bt);

    // ------------------------------------------------------------------ LM control (Spec a0)
    if (tid == 0) {
      // trial cost and invalid count: the camera pass at the trial state (already summed over the cluster)
      const double tc = step_ok ? scal[0] : 0.0, ninv = step_ok ? scal[1] : 0.0;
      double dDd = scal[2], gd = scal[3], dn2 = scal[4], xn2 = 0, gm = 0;
      if constexpr (CS > 1) {
        cg::cluster_group cl = cg::this_cluster();
        for (int r = 0; r < CS; ++r) {
          const double* xr = cl.map_shared_rank(xbuf, r) + Cfg::XDATA;
          dDd += xr[0]; gd += xr[1]; dn2 += xr[2]; xn2 += xr[3]; gm = fmax(gm, xr[4]);
        }
      } else {
        dDd += scal[9]; gd += scal[10]; dn2 += scal[11]; xn2 = scal[12]; gm = scal[13];
      }
      for (int i = 0; i < nc * 6; ++i) if (fidx[i / 6] >= 0) gm = fmax(gm, fabs(gc[i]));
      const double model = 0.5 * (dDd * imu - gd);
      int accept = 0, stop = 0;
      double rho = 0.0;
      if (step_ok && model > 0.0) {
        rho = (cost - tc) / model;
        // a step that pushes MORE observations behind the cheirality gate (they cost 0) is rejected like a failed one
        accept = (isfinite(tc) && rho > kMinRatio && (int)ninv <= ninv_cur) ? 1 : 0;
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
    // (cluster: the step scalars in the peers' xbufs are next overwritten after two more cluster barriers, so a CTA
    // barrier is enough here; one cluster barrier after the loop keeps every CTA's shared memory alive until all
    // remote reads are done)
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
      for (int i = tid; i < nc * PS; i += NTHREADS) pose_cur[i] = pose_tri[i];
      // the trial evaluation is the linearisation of the accepted state
      for (int i = tid; i < MAXCAM * 27; i += NTHREADS) Bc[i] = Bt[i];
      for (int i = tid; i < nc * 6; i += NTHREADS) Dc[i] = clampd(Bt[(i / 6) * 21 + packed21(i % 6, i % 6)]);
      cur ^= 1;
      fresh_lin = 1;
      ++n_acc;
    } else {
      mu = mu / nu;
      nu *= 2.0;
      fresh_lin = 0;
      ++n_rej;
    }
    if (tid == 0 && crank == 0 && d.hist_cost != nullptr && it < kHist) {
      d.hist_cost[(size_t)win * kHist + it] = cost;
      d.hist_mu[(size_t)win * kHist + it] = mu_used;
      d.hist_accept[(size_t)win * kHist + it] = accept;
    }
    if (!stop && !(mu > kMuMin)) stop = B2BA_NUMERICAL_FAILURE;
    __syncthreads();
    if (stop) { status = stop; ++it; break; }
  }

  // ------------------------------------------------------------------ write back
  if constexpr (CS > 1) cg::this_cluster().sync();   // no CTA leaves while a peer may still read its shared memory
  if (cur == 1) {
    const double* src = d.pts[1] + (size_t)p0 * 3;
    double* dst = d.pts[0] + (size_t)p0 * 3;
    for (int i = crank * NTHREADS + tid; i < npw * 3; i += CS * NTHREADS) dst[i] = src[i];
  }
  if (crank != 0) return;                  // poses and the summary are identical in every CTA of the cluster
  for (int i = tid; i < nc * 12; i += NTHREADS) d.poses[(size_t)c0 * 12 + i] = pose_cur[(i / 12) * PS + i % 12];
  if (tid == 0) {
    b2ba_summary s;
    s.status = status; s.iterations = it; s.accepted = n_acc; s.rejected = n_rej;
    s.n_invalid = ninv_cur; s.pcg_iters = 0;
    s.initial_cost = initial_cost; s.final_cost = cost; s.final_mu = mu;
    d.summaries[win] = s;
  }
}

using CfgSmall = WinCfg<7, 8, 4, 64, kSmallMaxCam, 2>;    // 256 threads, 2 CTAs / SM
#ifndef WX_G
#define WX_G 4
#endif
#ifndef WX_MINB
#define WX_MINB 4
#endif
using CfgSmall4 = WinCfg<7, 4, WX_G, 32, kSmallMaxCam, WX_MINB>;   // 128 threads, 4 CTAs / SM
using CfgBig = WinCfg<16, 16, 16, 32, kMaxCam, 1>;        // n+1 <= 128: up to 21 free cameras (C3)
constexpr int kClusterSize = 8;                           // portable cluster size
using CfgSmallC = WinCfg<7, 8, 4, 64, kSmallMaxCam, 1, kClusterSize>;   // latency path: 8 CTAs x 256 threads per window
using CfgBigC = WinCfg<16, 16, 16, 32, kMaxCam, 1, kClusterSize>;      // latency path, up to 21 free cameras
using CfgSmallC16 = WinCfg<7, 8, 4, 64, kSmallMaxCam, 1, 16>;          // <= 8 windows: 16 CTAs (non-portable cluster size)
using CfgBigC16 = WinCfg<16, 16, 16, 32, kMaxCam, 1, 16>;

// cudaFuncSetAttribute applies to the CURRENT device's copy of the kernel: the opt-in (dynamic shared memory above
// 48 KB, non-portable cluster size) is done once per (instantiation, device), under a mutex -- handles on different
// devices of one process are independent (include/b2ba.h)
template <class Cfg, int LOSS>
static int configure_once() {
  static std::mutex mtx;
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  std::lock_guard<std::mutex> lock(mtx);
  if (dev >= 0 && dev < 64 && configured[dev]) return 0;
  e = cudaFuncSetAttribute(window_lm_kernel<Cfg, LOSS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM);
  if (e != cudaSuccess) return (int)e;
  if (Cfg::CS > 8) {
    e = cudaFuncSetAttribute(window_lm_kernel<Cfg, LOSS>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return (int)e;
  }
  if (dev >= 0 && dev < 64) configured[dev] = true;
  return 0;
}

template <class Cfg, int LOSS>
static int launch_l(const WinData& d, const Params& p, cudaStream_t s) {
  if (const int rc = configure_once<Cfg, LOSS>()) return rc;
  if constexpr (Cfg::CS > 1) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)d.nb * Cfg::CS, 1, 1);
    lc.blockDim = dim3(Cfg::NTHREADS, 1, 1);
    lc.dynamicSmemBytes = Cfg::SMEM;
    lc.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = Cfg::CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    lc.attrs = at; lc.numAttrs = 1;
    return (int)cudaLaunchKernelEx(&lc, window_lm_kernel<Cfg, LOSS>, d, p);
  } else {
    window_lm_kernel<Cfg, LOSS><<<d.nb, Cfg::NTHREADS, Cfg::SMEM, s>>>(d, p);
    return (int)cudaGetLastError();
  }
}
template <class Cfg>
static int launch(const WinData& d, const Params& p, cudaStream_t s) {
  return p.loss == B2BA_LOSS_HUBER ? launch_l<Cfg, 1>(d, p, s) : launch_l<Cfg, 0>(d, p, s);
}

// variant: kWinAuto, or a forced instantiation (options.flags bits 2 / 3; the parity tests run every one of them).
// B2BA_WIN_VARIANT=0 (developer knob) makes the 256-thread CTA the default of the batched small path.
static bool small256_default() {
  static const bool v = [] { const char* e = std::getenv("B2BA_WIN_VARIANT"); return e && std::atoi(e) == 0; }();
  return v;
}
int launch_window_small(const WinData& d, const Params& p, cudaStream_t s, int variant) {
  return (variant & kWinSmall256) || small256_default() ? launch<CfgSmall>(d, p, s) : launch<CfgSmall4>(d, p, s);
}
int launch_window_big(const WinData& d, const Params& p, cudaStream_t s) { return launch<CfgBig>(d, p, s); }
// B2BA_CLUSTER16=0 (developer knob) keeps the portable cluster size for every window count
static bool cluster16() {
  static const bool v = [] { const char* e = std::getenv("B2BA_CLUSTER16"); return !e || std::atoi(e) != 0; }();
  return v;
}
int launch_window_small_cluster(const WinData& d, const Params& p, cudaStream_t s, int variant) {
  return (d.nb <= 8 && cluster16() && !(variant & kWinCluster8)) ? launch<CfgSmallC16>(d, p, s) : launch<CfgSmallC>(d, p, s);
}
int launch_window_big_cluster(const WinData& d, const Params& p, cudaStream_t s, int variant) {
  return (d.nb <= 8 && cluster16() && !(variant & kWinCluster8)) ? launch<CfgBigC16>(d, p, s) : launch<CfgBigC>(d, p, s);
}

template <class Cfg>
static int attrs_of(int* regs, int* smem_bytes, int* ctas_per_sm) {
  cudaFuncAttributes a;
  int nb = 0;
  cudaError_t e = cudaFuncGetAttributes(&a, window_lm_kernel<Cfg, 0>);
  if (e == cudaSuccess) {
    cudaFuncSetAttribute(window_lm_kernel<Cfg, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM);
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, window_lm_kernel<Cfg, 0>, Cfg::NTHREADS, Cfg::SMEM);
  }
  if (e != cudaSuccess) return (int)e;
  *regs = a.numRegs; *smem_bytes = (int)Cfg::SMEM; *ctas_per_sm = nb;
  return 0;
}

int window_kernel_attrs(int which, int* regs, int* smem_bytes, int* ctas_per_sm) {
  if (which == 0) return small256_default() ? attrs_of<CfgSmall>(regs, smem_bytes, ctas_per_sm)
                                            : attrs_of<CfgSmall4>(regs, smem_bytes, ctas_per_sm);
  return attrs_of<CfgBig>(regs, smem_bytes, ctas_per_sm);
}

}  // namespace b2
