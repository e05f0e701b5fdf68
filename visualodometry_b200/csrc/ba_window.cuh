// Window path (C1 / C3 / C4): ONE CTA runs the whole Levenberg-Marquardt loop of ONE window,
// all iteration state in shared memory / registers, zero host synchronisation per iteration.
// See DESIGN.md "Window kernel".  Kernel body: ba_window.cu.
#pragma once
#include "ba_math.cuh"
#include "../../include/b2ba.h"

namespace b2 {

constexpr int kMaxCam = 24;       // cameras per window on this path (free + fixed)
constexpr int kPtScratch = 12;    // per point: F(6) gp(3) D(3)
constexpr int kHist = 64;

struct WinData {
  int nb;
  const int* cam_off;          // nb+1
  const int* pt_off;           // nb+1
  const ObsRec* pm;            // point-major stream (idx = local camera)
  const int* pt_ptr;           // NpTot+1, absolute offsets into pm
  const int* sched;            // NpTot: per window, local landmark ids by decreasing observation count
  const ObsRec* cm;            // camera-major stream (idx = local point)
  const int* cam_ptr;          // NcTot+1, absolute offsets into cm
  const unsigned char* fixed;  // NcTot
  double* poses;               // NcTot x 12 (in/out)
  double* pts[2];              // NpTot x 3 ping-pong; pts[0] holds the result at exit
  double* pt_scratch;          // NpTot x kPtScratch
  b2ba_summary* summaries;     // nb
  double* hist_cost;           // nb x kHist or null
  double* hist_mu;
  int* hist_accept;
  double* dbg;                 // single-window stage dump or null (layout: WinDbg)
};

// layout of the stage dump (doubles), n = 6*nfree
struct WinDbg {
  int nc, np, n;
  __host__ __device__ size_t off_B() const { return 0; }
  __host__ __device__ size_t off_gc() const { return off_B() + (size_t)nc * 36; }
  __host__ __device__ size_t off_C() const { return off_gc() + (size_t)nc * 6; }
  __host__ __device__ size_t off_gp() const { return off_C() + (size_t)np * 6; }
  __host__ __device__ size_t off_S() const { return off_gp() + (size_t)np * 3; }
  __host__ __device__ size_t off_gred() const { return off_S() + (size_t)n * n; }
  __host__ __device__ size_t off_dc() const { return off_gred() + n; }
  __host__ __device__ size_t off_dp() const { return off_dc() + n; }
  __host__ __device__ size_t off_scal() const { return off_dp() + (size_t)np * 3; }
  __host__ __device__ size_t total() const { return off_scal() + 4; }
};

// Upper-triangle tile enumeration (row-major over NT tile rows), usable in constant expressions.
template <int NT>
struct Tri {
  static constexpr int T = NT * (NT + 1) / 2;
  static constexpr __host__ __device__ int mi(int t) { int m = 0, len = NT; while (t >= len) { t -= len; --len; ++m; } return m; }
  static constexpr __host__ __device__ int ni(int t) { int m = 0, len = NT; while (t >= len) { t -= len; --len; ++m; } return m + t; }
};

// NT tile rows (8 each) of the reduced system incl. the y row; NWARPS = G tile groups x KS k-splits;
// PB landmarks per batch; MAXCAM cameras per window held in shared memory; MINB CTAs per SM; CS = CTAs
// (one thread-block cluster) that share ONE window: landmarks / observations are split across the
// cluster, partial sums are all-reduced through distributed shared memory in a fixed order, and the
// reduced camera system is solved redundantly by every CTA (CS = 1: one CTA per window, no cluster).
template <int NT_, int NWARPS_, int G_, int PB_, int MAXCAM_, int MINB_, int CS_ = 1>
struct WinCfg {
  static constexpr int NT = NT_, NWARPS = NWARPS_, G = G_, PB = PB_, MAXCAM = MAXCAM_, MINB = MINB_, CS = CS_;
  static constexpr int NTHREADS = NWARPS * 32;
#ifndef WX_TILEPERM
#define WX_TILEPERM 1
#endif
#ifndef WX_POSE_STRIDE
#define WX_POSE_STRIDE 14
#endif
  // poses in shared memory: 12 doubles padded to 14, so that the LDS.128 of lanes holding up to 8 different cameras
  // fall into different banks (stride 12: cameras c and c + 4 collide)
  static constexpr int PS = WX_POSE_STRIDE;
  static constexpr int KS = NWARPS / G;            // k-splits of the SYRK
  static constexpr int T = NT * (NT + 1) / 2;      // upper 8x8 tiles
  static constexpr int TPW = (T + G - 1) / G;      // tiles per warp
  static constexpr int TPP = NTHREADS / PB;        // lanes per point
  // canonical tile (upper triangle, row-major) computed by warp group `grp` as its q-th tile.  For the 7 x 7 / 4-group
  // shape of the batched small path the partition is chosen so that a group touches few distinct tile rows (17 row
  // fragments per k-step over the 4 groups instead of 22 with consecutive tiles): a fragment is loaded from shared
  // memory once per group and k-step, and the shared-memory pipe is one of the two resources that bound the kernel.
  static constexpr __host__ __device__ int tile_of(int grp, int q) {
    if (NT == 7 && G == 4 && WX_TILEPERM) {
      // groups: rows {0,1,4,5,6}, {1,2,3,4}, {0,2,3,6}, {2,3,4,5}
      constexpr int tab[4][7] = {{1, 4, 5, 11, 12, 24, 26}, {7, 8, 9, 10, 13, 14, 15},
                                 {0, 2, 3, 6, 17, 21, 27}, {16, 18, 19, 20, 22, 23, 25}};
      return tab[grp][q];
    }
    return grp * TPW + q;
  }
  static constexpr int NR = 8 * NT;                // rows of Z (6*nfree + 1, padded)
  static constexpr int NCOL = 3 * PB;              // columns of Z per batch (3 per landmark)
  static constexpr int LDT = NR + 4;               // Z is stored TRANSPOSED: Zt[col * LDT + row];
                                                   // LDT == 4 or 12 mod 16 -> conflict-free DMMA fragment loads
  static constexpr int KSTEPS = NCOL / 4;          // k-steps (of 4 columns) per batch
  static constexpr int LDS = NR + 1;               // leading dim of S (odd)
  // the tile memory is reused after the point pass: [0, SPART) this CTA's Z Z^T tiles (T x 64, as the
  // DMMA accumulators hold them), [SPART, SPART + NR*LDS) the assembled system S
  static constexpr int SPART = T * 64;
  static constexpr int ZS = (NCOL * LDT > SPART + NR * LDS) ? NCOL * LDT : SPART + NR * LDS;
  static constexpr int NRED = NWARPS * 32;
  // cluster exchange buffer: [0, XDATA) camera blocks or this CTA's share of the reduced Z Z^T tiles,
  // [XDATA, XDATA + 8) step scalars, [XDATA + 8] step-bad flag
  static constexpr int XCHUNK = (T * 64 + (CS > 1 ? CS : 1) - 1) / (CS > 1 ? CS : 1);
  static constexpr int XDATA = ((MAXCAM * 27 + 2 > XCHUNK ? MAXCAM * 27 + 2 : XCHUNK) + 1) / 2 * 2;
  static constexpr int XBUF = CS > 1 ? XDATA + 16 : 0;
  // doubles: pose_cur, pose_tri, B(21), gc, Dc, dvec, gred, red, scal, xbuf, Zt/S
  static constexpr int SM_DOUBLES = MAXCAM * (PS + PS + 21 + 6 + 6) + 128 + 128 + NRED + 16 + XBUF + ZS;
  static constexpr size_t SMEM = (size_t)SM_DOUBLES * 8 + (2 * MAXCAM + 8) * 4;
  static_assert(KS * G == NWARPS, "warps = groups x k-splits");
  static_assert(KSTEPS % KS == 0, "k-steps split evenly");
  static_assert(TPP >= 1 && TPP <= 32 && (TPP & (TPP - 1)) == 0, "lanes per point: power of two");
  static_assert(NCOL % 4 == 0 && (MAXCAM % 2) == 0, "alignment");
  static_assert(6 * (MAXCAM - 1) + 1 <= NR || true, "n+1 <= NR is checked at path selection");
};

constexpr int kSmallMaxCam = 12, kSmallMaxFree = 9;    // window_small: n+1 <= 56
constexpr int kBigMaxFree = 21;                        // window_big:   n+1 <= 128

// launchers (ba_window.cu). Return cudaError_t as int.  `variant`: kWinAuto or forced instantiations (options.flags
// bits 2 / 3), so that every instantiation can be reached -- and parity-tested -- through the C ABI.
enum { kWinAuto = 0, kWinSmall256 = 1 /* batched small path: 256-thread CTAs, 2 per SM */,
       kWinCluster8 = 2 /* latency path: the portable 8-CTA cluster even for <= 8 windows */ };
int launch_window_small(const WinData& d, const Params& p, cudaStream_t s, int variant = kWinAuto);   // <= 12 cameras, <= 9 free (n+1 <= 56)
int launch_window_big(const WinData& d, const Params& p, cudaStream_t s);     // n+1 <= 128 (<= 21 free cams)
// Few windows (latency path): one 8- or 16-CTA cluster per window.  Same arithmetic per observation; sums are formed
// per CTA and then across the cluster in rank order, so results differ from the 1-CTA kernels by rounding only.
int launch_window_small_cluster(const WinData& d, const Params& p, cudaStream_t s, int variant = kWinAuto);
int launch_window_big_cluster(const WinData& d, const Params& p, cudaStream_t s, int variant = kWinAuto);
int window_kernel_attrs(int which, int* regs, int* smem_bytes, int* ctas_per_sm);

}  // namespace b2
