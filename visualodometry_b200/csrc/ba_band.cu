// Exact solver of the large path for BLOCK-BANDED reduced camera systems (a5/a6): when the
// cameras that share landmarks are close in index (VO / chain-like BAL visibility: C5 `local`), S is
// block-banded and a banded Cholesky costs n * (6 w)^2 flops instead of PCG iterations that do not converge
// on long chains (block-Jacobi PCG needs O(chain length) iterations).
//
// A banded factorisation is a chain of N dependent 6x6 pivots, i.e. latency-bound.  Two measures:
//  * per block column: one CTA of 128 threads (one warp per scheduler), the active window in a shared-memory
//    ring filled by cp.async three columns ahead, the pivot block factored redundantly in registers, the
//    forward substitution of all right-hand sides fused into the factorisation;
//  * across block columns: SUBSTRUCTURING.  The N block columns are cut into P interiors separated by P-1
//    separators of w block columns.  Interiors do not couple with each other, so P CTAs factor them
//    concurrently, each carrying [b | C_left | C_right] (its couplings to the neighbouring separators) as
//    1 + 12 w right-hand sides: Y = L^-1 [b | C].  The separator system  T = A_SS - sum_i Y_C^T Y_C  is again
//    block-banded ((P-1) w block columns, half-bandwidth 2w-1) and is solved by the same kernel with one CTA;
//    the interiors then back-substitute concurrently with  z = y_b - Y_C x_S.
// The chain shrinks from N to about N/P + (P-1) w columns (C5: 1000 -> ~120).  Deterministic (no atomics).
#include "ba_large.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>

namespace b2 {
namespace {

constexpr int kBandThreads = 128;
constexpr int kBandAhead = 3;      // block columns loaded ahead of the active window
constexpr int kBackRing = 8;       // backward pass: columns staged ahead in shared memory

// One SPD block-banded system.  band block (j, k), k = 0..w, is A_{j+k, j} (6 x 6 row-major): the lower
// band by block column (in: A, out: L).  rhs: per block column 6 x nr row-major (in: B, out: Y = L^-1 B, and
// after the backward pass column 0 holds z).  binv: 1 / diag(L).
struct BandSys {
  double* band;
  double* rhs;
  double* binv;
  int N, w, nr;
};

// interior i of P over N block columns with separators of w columns between them
__host__ __device__ inline void chunk_range(int N, int w, int P, int i, int& lo, int& hi) {
  const int M = N - (P - 1) * w, q = M / P, rem = M % P;
  lo = i * q + (i < rem ? i : rem) + i * w;
  hi = lo + q + (i < rem ? 1 : 0);
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// ------------------------------------------------------------------------------------------------ fill
// S (upper block CSR) -> lower band by column; right-hand sides of the P interiors ([b | C_left | C_right])
// or, for P == 1, the single right-hand side b = -g_red.
__global__ void band_fill_kernel(LargeData d, BandSys s, int P) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;
  const int W1 = s.w + 1;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  const size_t total = (size_t)s.N * W1 * 36;
  for (size_t i = gid; i < total; i += gsz) {
    const int e = (int)(i % 36), k = (int)((i / 36) % W1), j = (int)(i / (36 * (size_t)W1));
    const int r = e / 6, c = e % 6;
    double v = 0.0;
    // find block (j, j+k) in row j of the upper structure (rows are short: linear scan)
    for (int b = d.blk_ptr[j]; b < d.blk_ptr[j + 1]; ++b)
      if (d.blk_col[b] == j + k) {
        // diagonal block: only its upper triangle is maintained by the Schur pass
        v = (k == 0 && r < c) ? d.Sg[(size_t)b * 36 + 6 * r + c] : d.Sg[(size_t)b * 36 + 6 * c + r];
        break;
      }
    s.band[i] = v;
  }
}

// value of A_{row, col} [r][c] (block indices, any order) read from S; 0 outside the structure
__device__ inline double s_entry(const LargeData& d, int row, int col, int r, int c) {
  int a = row, b = col, rr = r, cc = c;
  if (a > b) { a = col; b = row; rr = c; cc = r; }       // S stores (a <= b); A_{b,a} = A_{a,b}^T
  for (int q = d.blk_ptr[a]; q < d.blk_ptr[a + 1]; ++q)
    if (d.blk_col[q] == b) {
      if (a == b && rr > cc) { const int t = rr; rr = cc; cc = t; }
      return d.Sg[(size_t)q * 36 + 6 * rr + cc];
    }
  return 0.0;
}

__global__ void band_fill_rhs_kernel(LargeData d, BandSys s, int P) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;
  const double* gred = d.Sg + (size_t)d.nblk * 36;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  const int nr = s.nr, w = s.w;
  const size_t total = (size_t)s.N * 6 * nr;
  for (size_t i = gid; i < total; i += gsz) {
    const int q = (int)(i % nr), r = (int)((i / nr) % 6), j = (int)(i / ((size_t)6 * nr));
    double v = 0.0;
    if (q == 0) {
      v = -gred[6 * j + r];
    } else {
      // which interior holds column j?  (separator columns keep zero rows: they are never read)
      int ci = -1, lo = 0, hi = 0;
      for (int t = 0; t < P; ++t) {
        chunk_range(s.N, w, P, t, lo, hi);
        if (j >= lo && j < hi) { ci = t; break; }
        if (j < lo) break;
      }
      if (ci >= 0) {
        const int qq = q - 1, side = qq / (6 * w), a = (qq % (6 * w)) / 6, c = qq % 6;
        const int sep = side == 0 ? lo - w + a : hi + a;          // separator block column
        const bool exists = side == 0 ? ci > 0 : ci < P - 1;
        if (exists && abs(sep - j) <= w) v = s_entry(d, j, sep, r, c);
      }
    }
    s.rhs[i] = v;
  }
}

// ------------------------------------------------------------------------------------------------ forward
// CTA i factors the block columns [lo_i, hi_i) of `s` and forward-substitutes their nr right-hand sides.
template <int NT>
__global__ void __launch_bounds__(NT) band_forward_kernel(LargeCtrl* ctrl, BandSys s, int P) {
  if (ctrl->stop || ctrl->step_bad) return;
  extern __shared__ __align__(16) double bsm[];
  const int tid = threadIdx.x;
  const int w = s.w, W1 = w + 1, nr = s.nr;
  int lo, hi;
  chunk_range(s.N, w, P, blockIdx.x, lo, hi);
  const int NB = W1 * 36;                              // band part of a slot
  const int SLOT = NB + 6 * nr;                        // one block column (w+1 blocks) + its right-hand sides
  const int RS = W1 + kBandAhead;                      // ring slots: the active window + columns loaded ahead
  double* ring = bsm;
  double* Lsm = ring + (size_t)RS * SLOT;              // factor of the current diagonal block (36)
  __shared__ int bad;
  __shared__ unsigned char pairA[kBandMaxBw * (kBandMaxBw + 1) / 2], pairB[kBandMaxBw * (kBandMaxBw + 1) / 2];
  if (tid == 0) {
    bad = 0;
    int k = 0;                                         // block pairs (a >= b >= 1) of a full window, b-major
    for (int b = 1; b <= w; ++b)
      for (int a = b; a <= w; ++a) { pairA[k] = (unsigned char)a; pairB[k] = (unsigned char)b; ++k; }
  }
  const int npair_full = w * (w + 1) / 2;

  auto load_col = [&](int j, int slot) {               // global band column j + its rhs rows -> ring slot
    double* sl = ring + (size_t)slot * SLOT;
    const double* g = s.band + (size_t)j * NB;
    for (int i = tid; i < NB / 2; i += NT) cp_async16(sl + 2 * i, g + 2 * i);
    const double* gr = s.rhs + (size_t)j * 6 * nr;
    for (int i = tid; i < 3 * nr; i += NT) cp_async16(sl + NB + 2 * i, gr + 2 * i);
  };
  for (int j = lo; j < min(lo + RS - 1, hi); ++j) load_col(j, j - lo);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();

  int sj = 0;                                          // ring slot of column j (incremental: no division)
  for (int j = lo; j < hi; ++j, sj = (sj + 1 == RS ? 0 : sj + 1)) {
    double* cj = ring + (size_t)sj * SLOT;
    auto slot_of = [&](int a) { const int t = sj + a; return ring + (size_t)(t >= RS ? t - RS : t) * SLOT; };
    // ---- every thread factors the diagonal block redundantly in registers
    double L[6][6], inv[6];
    bool isbad = false;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c)
        if (c <= r) L[r][c] = cj[6 * r + c];
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      double dj = L[c][c];
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (k < c) dj = fma(-L[c][k], L[c][k], dj);
      if (!(dj > 0.0) || !isfinite(dj)) { isbad = true; dj = 1.0; }
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
    if (isbad && tid == 0) bad = 1;
    const int wj = min(w, hi - 1 - j);                 // blocks below the diagonal inside this interior
    // ---- panel rows L_{j+k,j} = A_{j+k,j} L_jj^-T (one item per row) and Y_j = L_jj^-1 B_j (one item per
    // right-hand side)
    for (int it = tid; it < 6 * wj + nr; it += NT) {
      if (it < 6 * wj) {
        double* row = cj + 36 + 6 * it;                // row (it % 6) of block 1 + it / 6
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
      } else {
        double* b = cj + NB + (it - 6 * wj);           // column of the 6 x nr right-hand-side block
        double y[6];
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          double v = b[c * nr];
#pragma unroll
          for (int k = 0; k < 6; ++k)
            if (k < c) v = fma(-L[c][k], y[k], v);
          y[c] = v * inv[c];
        }
#pragma unroll
        for (int c = 0; c < 6; ++c) b[c * nr] = y[c];
      }
    }
    if (tid == NT - 1) {
#pragma unroll
      for (int r = 0; r < 6; ++r) {
#pragma unroll
        for (int c = 0; c < 6; ++c) Lsm[6 * r + c] = c <= r ? L[r][c] : 0.0;   // (block 0 of cj is still being read)
        s.binv[6 * (size_t)j + r] = inv[r];
      }
    }
    cp_async_wait<kBandAhead - 1>();                   // the column that enters the window at this step has landed
    __syncthreads();
    // ---- trailing update inside the window: A_{j+a,j+b} -= L_{j+a,j} L_{j+b,j}^T (a >= b >= 1), items = 6x6
    // entries flattened over (b, a, r, c); right-hand sides B_{j+a} -= L_{j+a,j} Y_j, items = (r, column)
    for (int id = tid; id < npair_full * 36; id += NT) {
      const int e = id % 36, pr = id / 36;
      const int a = pairA[pr], b = pairB[pr], r = e / 6, c = e % 6;
      if (a <= wj) {
        const double* la = cj + 36 * a + 6 * r;
        const double* lb = cj + 36 * b + 6 * c;
        double* dst = slot_of(b) + 36 * (a - b) + 6 * r + c;
        double v = *dst;
#pragma unroll
        for (int m = 0; m < 6; ++m) v = fma(-la[m], lb[m], v);
        *dst = v;
      }
    }
    // items = (block a, right-hand side q): the 6 rows of B_{j+a}[:, q] with Y_j[:, q] in registers
    for (int id = tid; id < wj * nr; id += NT) {
      int q = id, a = 1;
      while (q >= nr) { q -= nr; ++a; }
      double y[6];
#pragma unroll
      for (int m = 0; m < 6; ++m) y[m] = cj[NB + m * nr + q];
      const double* la = cj + 36 * a;
      double* dst = slot_of(a) + NB + q;
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        double v = dst[r * nr];
#pragma unroll
        for (int m = 0; m < 6; ++m) v = fma(-la[6 * r + m], y[m], v);
        dst[r * nr] = v;
      }
    }
    // ---- column j is final: L blocks back to the band array, Y_j to the rhs array
    {
      double* g = s.band + (size_t)j * NB;
      for (int i = tid; i < NB; i += NT) g[i] = i < 36 ? Lsm[i] : cj[i];
      double* gr = s.rhs + (size_t)j * 6 * nr;
      for (int i = tid; i < 6 * nr; i += NT) gr[i] = cj[NB + i];
    }
    // the slot of column j-1 is free since the previous barrier: it takes column j-1+RS, which is first
    // touched kBandAhead steps from now (asynchronous copy: nobody waits for it here)
    if (j + RS - 1 < hi) load_col(j + RS - 1, sj == 0 ? RS - 1 : sj - 1);
    cp_async_commit();
    __syncthreads();
  }
  cp_async_wait<0>();
  if (bad && tid == 0) ctrl->step_bad = 1;
}

// ------------------------------------------------------------------------------------------------ Gram
// G_i = Y_i^T Y_i (nr x nr) over the rows of interior i.  grid (P, ceil(nr / 4)): CTA (i, g) computes the
// rows 4g..4g+3 of G_i; thread (slice, q): column q, the rows of Y congruent to `slice`; the slices are
// added in a fixed order through shared memory.
constexpr int kGramThreads = 256;
__global__ void __launch_bounds__(kGramThreads) band_gram_kernel(LargeCtrl* ctrl, BandSys s, int P, double* gram) {
  if (ctrl->stop || ctrl->step_bad) return;
  __shared__ double red[kGramThreads * 4];
  const int nr = s.nr, CQ = nr <= 64 ? 64 : 128, nsl = kGramThreads / CQ;
  const int q = threadIdx.x % CQ, sl = threadIdx.x / CQ, p0 = 4 * blockIdx.y;
  int lo, hi;
  chunk_range(s.N, s.w, P, blockIdx.x, lo, hi);
  double acc[4] = {0, 0, 0, 0};
  const int rows = 6 * (hi - lo);
  int pi[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) pi[k] = min(p0 + k, nr - 1);
  if (q < nr) {
    const double* y = s.rhs + ((size_t)lo * 6 + sl) * nr;
#pragma unroll 4
    for (int r = sl; r < rows; r += nsl, y += (size_t)nsl * nr) {
      const double yq = y[q];
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[k] = fma(y[pi[k]], yq, acc[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) red[(sl * 4 + k) * CQ + q] = acc[k];
  __syncthreads();
  if (sl == 0 && q < nr) {
    double* g = gram + (size_t)blockIdx.x * nr * nr;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      double v = 0.0;
      for (int t = 0; t < nsl; ++t) v += red[(t * 4 + k) * CQ + q];
      if (p0 + k < nr) g[(size_t)(p0 + k) * nr + q] = v;
    }
  }
}

// ------------------------------------------------------------------------------------------------ separators
// The separator system T x_S = b_S' as a band system `t` (N' = (P-1) w block columns, half-bandwidth 2w-1,
// one right-hand side).  Separator i sits between interiors i and i+1: it is the RIGHT separator of interior
// i (Gram columns 1 + 6w + 6a + c) and the LEFT separator of interior i+1 (Gram columns 1 + 6a + c).
__global__ void band_sep_assemble_kernel(LargeData d, BandSys s, int P, const double* gram, BandSys t) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;
  const double* gred = d.Sg + (size_t)d.nblk * 36;
  const int w = s.w, nr = s.nr, W1t = t.w + 1;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  const size_t nband = (size_t)t.N * W1t * 36, total = nband + (size_t)t.N * 6;
  for (size_t i = gid; i < total; i += gsz) {
    if (i < nband) {
      const int e = (int)(i % 36), k = (int)((i / 36) % W1t), jc = (int)(i / (36 * (size_t)W1t));
      const int r = e / 6, c = e % 6;                  // entry [r][c] of block (row jc + k, column jc)
      const int si = jc / w, a = jc % w;               // column: separator si, local block a
      const int jr = jc + k, sj = jr / w, b = jr % w;  // row: separator sj, local block b
      double v = 0.0;
      if (jr < t.N) {
        int lo, hi;
        if (sj == si) {
          chunk_range(s.N, w, P, si, lo, hi);          // separator si = block columns hi .. hi + w - 1
          v = s_entry(d, hi + b, hi + a, r, c);
          const double* gl = gram + (size_t)si * nr * nr;         // interior si: right separator
          const double* gn = gram + (size_t)(si + 1) * nr * nr;   // interior si+1: left separator
          v -= gl[(size_t)(1 + 6 * w + 6 * b + r) * nr + (1 + 6 * w + 6 * a + c)];
          v -= gn[(size_t)(1 + 6 * b + r) * nr + (1 + 6 * a + c)];
        } else if (sj == si + 1) {
          const double* gn = gram + (size_t)sj * nr * nr;         // interior sj couples separators sj-1 and sj
          v = -gn[(size_t)(1 + 6 * w + 6 * b + r) * nr + (1 + 6 * a + c)];
        }
      }
      t.band[i] = v;
    } else {
      const size_t q = i - nband;
      const int jc = (int)(q / 6), r = (int)(q % 6), si = jc / w, a = jc % w;
      int lo, hi;
      chunk_range(s.N, w, P, si, lo, hi);
      const double* gl = gram + (size_t)si * nr * nr;
      const double* gn = gram + (size_t)(si + 1) * nr * nr;
      t.rhs[q] = -gred[6 * (hi + a) + r] - gl[(size_t)(1 + 6 * w + 6 * a + r) * nr] - gn[(size_t)(1 + 6 * a + r) * nr];
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward
// CTA i: z = y_b - Y_C x_S over its interior (all threads), then L^T x = z by warp 0 (x blocks of the last
// w columns stay in shared memory); x -> xout.  xsep (null for P == 1): the separator solution, (P-1) w x 6.
__global__ void __launch_bounds__(kBandThreads) band_backward_kernel(LargeCtrl* ctrl, BandSys s, int P,
                                                                     const double* xsep, double* xout) {
  if (ctrl->stop || ctrl->step_bad) return;
  extern __shared__ __align__(16) double bsm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int w = s.w, W1 = w + 1, nr = s.nr, NB = W1 * 36;
  int lo, hi;
  chunk_range(s.N, w, P, blockIdx.x, lo, hi);
  double* xs = bsm;                                    // ring of the last W1 solution blocks (6 W1)
  double* xc = xs + 6 * W1;                            // [x_left (6w) | x_right (6w)] of this interior
  if (xsep) {
    const int i = blockIdx.x;
    for (int k = tid; k < 12 * w; k += kBandThreads) {
      const int side = k / (6 * w), e = k % (6 * w);
      const int sep = side == 0 ? i - 1 : i;
      const double v = (sep >= 0 && sep < P - 1) ? xsep[(size_t)sep * 6 * w + e] : 0.0;
      xc[k] = v;
      // separator i's solution goes out with interior i
      if (side == 1 && sep < P - 1) xout[6 * (size_t)hi + e] = v;
    }
    __syncthreads();
    for (int row = tid; row < 6 * (hi - lo); row += kBandThreads) {
      double* y = s.rhs + ((size_t)lo * 6 + row) * nr;
      double z = y[0];
      for (int q = 1; q < nr; ++q) z = fma(-y[q], xc[q - 1], z);
      y[0] = z;
    }
    __syncthreads();
  }
  // Warp 0 walks the columns backwards.  Lane layout: c = lane % 6 (component), sl = lane / 6 (a-slice, 5
  // slices; lanes 30, 31 idle).  Column j's L blocks, 1/diag and z come from a shared-memory ring that the warp
  // itself fills kBackRing columns ahead with cp.async, so no global-memory latency sits on the serial chain.
  if (warp == 0) {
    const int c = lane % 6, sl = lane / 6;
    const int CS = NB + 12;                            // ring slot: L column | 1/diag (6) | z (6)
    double* bring = xc + 12 * w;
    auto issue = [&](int j, int slot) {
      double* dst = bring + (size_t)slot * CS;
      const double* g = s.band + (size_t)j * NB;
      for (int i = lane; i < NB / 2; i += 32) cp_async16(dst + 2 * i, g + 2 * i);
      if (lane < 3) cp_async16(dst + NB + 2 * lane, s.binv + 6 * (size_t)j + 2 * lane);
      else if (lane >= 8 && lane < 14) cp_async8(dst + NB + 6 + (lane - 8), s.rhs + ((size_t)j * 6 + (lane - 8)) * nr);
    };
    for (int k = 0; k < kBackRing; ++k) {
      if (hi - 1 - k >= lo) issue(hi - 1 - k, k);
      cp_async_commit();
    }
    int slot = 0, xj = (hi - 1) % W1;                   // ring slot of column j; x-ring slot j % W1
    for (int j = hi - 1; j >= lo; --j) {
      cp_async_wait<kBackRing - 1>();
      __syncwarp();
      const double* col = bring + (size_t)slot * CS;
      const int wj = min(w, hi - 1 - j);
      const double inv_c = lane < 6 ? col[NB + lane] : 0.0, yc = lane < 6 ? col[NB + 6 + lane] : 0.0;
      double dg[6];
#pragma unroll
      for (int r = 0; r < 6; ++r) dg[r] = lane < 6 ? col[6 * r + c] : 0.0;   // column c of L_jj: L[r][c], r > c used
      // t[c] = sum_a sum_r L_{j+a,j}[r][c] x_{j+a}[r]
      double t = 0.0;
      if (lane < 30) {
        for (int a = 1 + sl; a <= wj; a += 5) {
          int xa_i = xj + a;
          while (xa_i >= W1) xa_i -= W1;
          const double* bl = col + 36 * a + c;
          const double* xa = xs + 6 * xa_i;
#pragma unroll
          for (int r = 0; r < 6; ++r) t = fma(bl[6 * r], xa[r], t);
        }
      }
      t += __shfl_down_sync(0xffffffffu, t, 6) + __shfl_down_sync(0xffffffffu, t, 12) +
           __shfl_down_sync(0xffffffffu, t, 18) + __shfl_down_sync(0xffffffffu, t, 24);
      // lanes 0..5: solve L_jj^T x = z - t by back substitution, one shuffle per component
      double xv = yc - t;                               // lane c holds component c
#pragma unroll
      for (int i = 5; i >= 0; --i) {
        const double xi = __shfl_sync(0xffffffffu, xv * inv_c, i);   // x_i is final when its turn comes
        if (lane == i) xv = xi;
        else if (lane < i) xv = fma(-dg[i], xi, xv);    // L[i][c] of lane c = dg[i]
      }
      if (lane < 6) { xs[6 * xj + lane] = xv; xout[6 * (size_t)j + lane] = xv; }
      __syncwarp();
      if (j - kBackRing >= lo) issue(j - kBackRing, slot);   // the slot just consumed takes the next column
      cp_async_commit();
      slot = slot + 1 == kBackRing ? 0 : slot + 1;
      xj = xj == 0 ? W1 - 1 : xj - 1;
    }
    cp_async_wait<0>();
  }
  if (blockIdx.x == 0 && tid == 0) ctrl->pcg_last = 0;
}

// ------------------------------------------------------------------------------------------------ host
struct BandPlan {
  int P, nr, Nt, wt;
  size_t off_rhs, off_gram, off_tband, off_trhs, off_tbinv, off_xsep, total;
};

inline size_t even(size_t x) { return (x + 1) & ~(size_t)1; }

BandPlan make_plan(int N, int w, int P) {
  BandPlan p{};
  p.P = P;
  p.nr = P > 1 ? 1 + 12 * w : 1;
  p.Nt = (P - 1) * w;
  p.wt = P > 1 ? 2 * w - 1 : 0;
  size_t o = 0;
  p.off_rhs = o;   o += even((size_t)N * 6 * p.nr);
  p.off_gram = o;  o += even((size_t)P * p.nr * p.nr);
  p.off_tband = o; o += even((size_t)p.Nt * (p.wt + 1) * 36);
  p.off_trhs = o;  o += even((size_t)p.Nt * 6);
  p.off_tbinv = o; o += even((size_t)p.Nt * 6);
  p.off_xsep = o;  o += even((size_t)p.Nt * 6);
  p.total = o;
  return p;
}

size_t fwd_smem(int w, int nr) {
  const size_t W1 = (size_t)w + 1;
  return ((W1 + kBandAhead) * (W1 * 36 + 6 * (size_t)nr) + 36) * sizeof(double);
}
size_t bwd_smem(int w) { return ((size_t)6 * (w + 1) + 12 * (size_t)w + (size_t)kBackRing * ((size_t)(w + 1) * 36 + 12)) * sizeof(double); }

}  // namespace

// number of interiors for N block columns of half-bandwidth w (1 = plain sequential factorisation)
int large_band_chunks(int N, int w) {
  if (const char* e = getenv("B2BA_BAND_CHUNKS")) {
    const int P = atoi(e);
    if (P >= 1 && (P == 1 || (w >= 1 && w <= kBandParMaxBw && (N - (P - 1) * w) / P >= w))) return P;
    return 1;
  }
  if (w < 1 || w > kBandParMaxBw) return 1;
  // chain length ~ N/P + 1.3 (P-1) w (the separator system has the wider band)
  int P = (int)std::sqrt((double)N / (1.3 * w));
  while (P > 1 && (N - (P - 1) * w) / P < 2 * w) --P;
  return P < 2 ? 1 : P;
}

size_t large_band_ws_count(int N, int w, int P) { return make_plan(N, w, P).total; }

size_t large_band_smem(int bw) { return fwd_smem(bw, 1); }

int large_band_solve(const LargeData& d, cudaStream_t st) {
  static size_t configured = 0;
  const int N = d.nfree, w = d.bw, P = d.band_P;
  const BandPlan pl = make_plan(N, w, P);
  const size_t smem = std::max(fwd_smem(w, pl.nr), P > 1 ? fwd_smem(pl.wt, 1) : (size_t)0);
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(band_forward_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(band_forward_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(band_forward_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    configured = smem;
  }
  static size_t configured_b = 48 * 1024;
  const size_t smem_b = std::max(bwd_smem(w), P > 1 ? bwd_smem(pl.wt) : (size_t)0);
  if (smem_b > configured_b) {
    cudaError_t e = cudaFuncSetAttribute(band_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b);
    if (e != cudaSuccess) return (int)e;
    configured_b = smem_b;
  }
  // wide bands: the trailing update is throughput-bound, 256 threads; narrow bands: latency-bound, one warp
  // per scheduler
  static const int env_threads = getenv("B2BA_BAND_THREADS") ? atoi(getenv("B2BA_BAND_THREADS")) : 0;
  auto forward = [&](const BandSys& sys, int chunks, size_t sm) {
    const int nt = env_threads ? env_threads : 256;
    if (nt >= 512) band_forward_kernel<512><<<chunks, 512, sm, st>>>(d.ctrl, sys, chunks);
    else if (nt >= 256) band_forward_kernel<256><<<chunks, 256, sm, st>>>(d.ctrl, sys, chunks);
    else band_forward_kernel<128><<<chunks, 128, sm, st>>>(d.ctrl, sys, chunks);
  };
  BandSys s{d.band, d.band_ws + pl.off_rhs, d.binv, N, w, pl.nr};
  band_fill_kernel<<<2 * d.sms, 256, 0, st>>>(d, s, P);
  band_fill_rhs_kernel<<<2 * d.sms, 256, 0, st>>>(d, s, P);
  forward(s, P, fwd_smem(w, pl.nr));
  if (P > 1) {
    double* gram = d.band_ws + pl.off_gram;
    BandSys t{d.band_ws + pl.off_tband, d.band_ws + pl.off_trhs, d.band_ws + pl.off_tbinv, pl.Nt, pl.wt, 1};
    double* xsep = d.band_ws + pl.off_xsep;
    band_gram_kernel<<<dim3(P, (pl.nr + 3) / 4), kGramThreads, 0, st>>>(d.ctrl, s, P, gram);
    band_sep_assemble_kernel<<<d.sms, 256, 0, st>>>(d, s, P, gram, t);
    forward(t, 1, fwd_smem(pl.wt, 1));
    band_backward_kernel<<<1, kBandThreads, bwd_smem(pl.wt), st>>>(d.ctrl, t, 1, nullptr, xsep);
    band_backward_kernel<<<P, kBandThreads, bwd_smem(w), st>>>(d.ctrl, s, P, xsep, d.dc);
  } else {
    band_backward_kernel<<<1, kBandThreads, bwd_smem(w), st>>>(d.ctrl, s, 1, nullptr, d.dc);
  }
  return (int)cudaGetLastError();
}

}  // namespace b2
