// Large path (C5, BAL-style; any camera count): multi-kernel Levenberg-Marquardt with a
// block-sparse reduced camera system and block-Jacobi PCG, all control on the device (the host
// only enqueues a fixed kernel sequence per LM iteration; no host synchronisation in the loop).
//
// Multi-GPU (SURVEY §8 row e), one process per GPU.  The small state (poses, landmarks, LM control) is REPLICATED and
// kept bit-identical on every rank; what is big is sharded and exchanged by the engine's own kernels through PEER
// MEMORY (every rank's exchange arena is mapped into every other rank with CUDA IPC; stores over NVLink / NVSwitch
// followed by a system-scope release of an epoch flag; no collective library call inside the LM loop):
//   * the reduced camera system S is sharded by BLOCK ROW: a rank forms the full rows it owns (Schur pair lists of
//     ALL landmarks, restricted to its rows) -- S is born sharded, it is never reduced or gathered for PCG;
//   * PCG runs on every rank in lockstep with replicated vectors; a rank multiplies its rows and stores its slice of
//     q = S p into every rank's q buffer: ONE all-gather of 6 doubles per owned row per CG iteration, inside the
//     kernel (dense C5: 18 MB of S per rank at 8 GPUs, L2-resident, instead of 144 MB streamed from HBM);
//   * banded / dense solvers: the owned rows are pushed into every rank's copy of S (small when S is banded) and the
//     factorisation runs replicated;
//   * back-substitution is sharded by landmark range; a rank stores its trial landmarks and its 8 partial sums into
//     every rank, which adds the partial sums in rank order (deterministic, identical everywhere).
// NCCL is only the bootstrap (unique id, exchange of the IPC handles, barriers at set-up).
#pragma once
#include "ba_math.cuh"
#include "devbuf.cuh"
#include "../../include/b2ba.h"

namespace b2 {

constexpr int kMaxRanks = 8;

// Peer-memory exchange state (nranks > 1).  Every rank allocates one arena with the SAME layout; arena[r] is rank r's
// arena as mapped into this process.  flags: kMaxRanks int64 per rank, flags[src] = last epoch rank `src` has
// published; epochs only grow (LargeCtrl::xepoch, identical on every rank because every rank runs the same exchanges).
struct Peer {
  double* arena[kMaxRanks];
  int rank, nranks;
  size_t off_flags, off_sums, off_qx, off_minv, off_r, off_sg, off_pts_tri;   // in doubles
  __host__ __device__ double* at(int r, size_t off) const { return arena[r] + off; }
};

struct LargeCtrl {          // device-resident LM state
  long long xepoch;         // peer exchanges completed so far (replicated)
  int xerr, xpad;           // a peer exchange timed out (the solve is abandoned, b2ba_solve reports it)
  double mu, nu, cost, initial_cost;
  double sums[8];           // reduced: trial cost, dDd, gd, dn2, xn2, n_invalid(trial), gmax, -
  double cam_sums[4];       // camera part of dDd, gd, dn2
  double cur_ninv;
  int it, n_acc, n_rej, status, stop, step_bad, have_lin, accept, pcg_iters, pcg_last, fresh_lin, pad;
};

struct LargeData {
  int nc, np, nfree, n;
  int p_lo, p_hi;                  // this rank's landmark range (whole problem on one GPU)
  int rank, nranks;
  int sms;                         // SM count of the handle's device (grid sizing)
  Peer peer;                       // nranks > 1: exchange arenas of all ranks
  int f_lo, f_hi;                  // nranks > 1: block rows of S this rank owns
  const int* lrow_ptr;             // (f_hi - f_lo + 1): the owned FULL rows (both triangles), local entry offsets
  const int* lcol;                 // column block of each local entry
  const int* lsrc;                 // upper block the entry is (or is the transpose of): id | (transposed << 31)
  const int* ldiag;                // (f_hi - f_lo): local entry index of each owned row's diagonal block
  int nlocal;                      // local entries
  double* Sl;                      // nlocal x 36: the owned rows of S, block (f, c) row-major
  const ObsRec* pm;                // point-major stream of the WHOLE problem (idx = camera)
  const int* pt_ptr;               // np+1
  const ObsRec* cm;                // camera-major stream of the LOCAL observations (idx = point)
  const int* cam_ptr;              // nc+1
  const int* fidx;                 // nc
  const int* blk_ptr;              // nfree+1   upper block structure of S (fi <= fj)
  const int* blk_col;
  int nblk;
  const int* frow_ptr;             // nfree+1   full symmetric structure for the mat-vec
  const int* fcol;
  const int* fsrc;                 // upper slot | (transposed << 31)
  double* poses;                   // nc x 12 current
  double* poses_tri;               // nc x 12 trial
  double* pts;                     // np x 3 current
  double* pts_tri;                 // np x 3 trial
  double* lin_part;                // local partial [nc*21 B | nc*6 g_c | cost | n_invalid]
  double* lin;                     // reduced (same layout)
  double* Dc;                      // nc x 6
  double* pt_scratch;              // np x 12 (F, g_p, D_p)
  double* Sg;                      // [nblk*36 S blocks | n g_red]   (one buffer: one allreduce)
  double* dc;                      // n   delta_c (PCG x)
  double* r; double* z; double* q; // n each
  double* Minv;                    // nfree x 36
  double* tbuf;                    // nblk x 6: PCG mat-vec, transposed products S(f,c)^T p_f of the upper blocks
  const int* pair_ptr;             // nblk+1: Schur pair list by S block (this rank's landmarks)
  const int2* pair_ab;             // (observation a, observation b >= a) in point-major numbering
  const int* blk_row;              // nblk: block row of each upper block
  const int* pm_pt;                // no: landmark of each point-major observation (this rank's range)
  int a_lo, a_hi;                  // this rank's observation range in the point-major stream
  int pair_lanes;                  // lanes per S block in the Schur gather (8 or 32)
  double* zbuf;                    // no x 24: per observation Z = E F (18) and Z F^T g_p (6)
  int bw;                          // block half-bandwidth of S if the banded Cholesky is used, else -1 (PCG)
  double* band;                    // nfree x (bw+1) x 36: lower band by block column (A, then L in place)
  double* binv;                    // nfree x 6: 1 / diag(L)
  int band_P;                      // interiors factored concurrently (1 = sequential), see ba_band.cu
  double* band_ws;                 // right-hand sides, Gram blocks and the separator system (large_band_ws_count)
  double* dense;                   // dense Cholesky (ba_dense.cu): (dense_ld + 64) x dense_ld lower matrix + rhs rows, else null
  int dense_ld;                    // n rounded up to the 64-column panel width
  double* dense_winv;              // dense_ld / 64 inverse diagonal factors (64 x 64 each)
  double* dense_x;                 // dense_ld solution blocks published by the backward pass
  int* dense_flags;                // dense_ld / 64 publication flags (epoch numbers)
  double* part;                    // partial-sum scratch (grid partials)
  double* sums_part;               // 8 local partial sums (before the allreduce)
  LargeCtrl* ctrl;
  b2ba_summary* summary;
  double* hist_cost; double* hist_mu; int* hist_accept;
  __host__ __device__ size_t lin_count() const { return (size_t)nc * 27 + 2; }
  __host__ __device__ size_t sg_count() const { return (size_t)nblk * 36 + n; }
};

// S block structure from the point-major stream (a8, device). Fills the DevBufs; returns cudaError.
struct LargeIndex {
  DevBuf<int> fidx, blk_ptr, blk_col, frow_ptr, fcol, fsrc, pair_ptr, blk_row, pm_pt;
  DevBuf<int> lrow_ptr, lcol, lsrc, ldiag;   // nranks > 1: the owned full rows (large_build_rows)
  int f_lo = 0, f_hi = 0, nlocal = 0;
  int a_lo = 0, a_hi = 0;
  DevBuf<int2> pair_ab;
  long long npairs = 0;
  int nblk = 0, nfull = 0, nfree = 0;
};
int large_build_structure(int nc, int np, long long no, const ObsRec* pm, const int* pt_ptr,
                          const unsigned char* h_fixed, LargeIndex& ix, cudaStream_t s);
// Schur pair list of the landmarks [p_lo, p_hi) (needs the structure): ix.pair_ptr / pair_ab / blk_row
int large_build_pairs(int p_lo, int p_hi, long long no, const ObsRec* pm, const int* pt_ptr, LargeIndex& ix, cudaStream_t s);
// nranks > 1: block rows [f_lo, f_hi) owned by `rank` (contiguous, balanced by full-row entries) and their local CSR
int large_build_rows(int rank, int nranks, LargeIndex& ix, cudaStream_t s);
// camera-major stream of the observations of landmarks [p_lo, p_hi)
int large_build_cm_slice(int nc, int p_lo, int p_hi, const ObsRec* pm, const int* pt_ptr, ObsRec* cm,
                         int* cam_ptr, cudaStream_t s);

int large_init_ctrl(const LargeData& d, const Params& P, cudaStream_t s);
// enqueue one LM iteration (nranks > 1: the peer-memory exchanges are part of the kernels)
// ev: null, or 7 events recorded at the phase boundaries (include/b2ba.h B2BA_T_*)
int large_iteration(const LargeData& d, const Params& P, int pcg_max_iters, double pcg_tol, cudaStream_t s, cudaEvent_t* ev = nullptr);
int large_pcg_dist_grid(int n, int nranks, int* blocks, size_t* smem);   // cooperative launch geometry of the distributed PCG
int large_finish(const LargeData& d, cudaStream_t s);
int large_pcg_grid(int n, int* blocks, size_t* smem);   // cooperative launch geometry
// exact banded Cholesky solve of S dc = -g_red (ba_band.cu); needs d.bw >= 0
size_t large_band_smem(int bw);
int large_band_chunks(int N, int bw);                 // interiors factored concurrently (1 = sequential)
size_t large_band_ws_count(int N, int bw, int P);     // doubles of workspace
int large_band_solve(const LargeData& d, cudaStream_t s);
// exact dense blocked Cholesky with DMMA trailing updates (ba_dense.cu); needs d.dense
int large_dense_npad(int n);
size_t large_dense_count(int n);                      // doubles of matrix storage
int large_dense_max_n();                              // one resident CTA per 64-block in the backward pass
int large_dense_solve(const LargeData& d, cudaStream_t s);
constexpr int kBandMaxBw = 22;     // (bw+4) x (bw+1) blocks of shared memory: 26 x 23 x 288 B = 172 KB
constexpr int kBandParMaxBw = 6;   // substructured solve: 1 + 12 bw right-hand sides per block column in the ring

}  // namespace b2
