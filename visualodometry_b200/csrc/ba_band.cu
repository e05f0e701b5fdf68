// Exact solver of the large path for BLOCK-BANDED reduced camera systems (a5/a6): when the
// cameras that share landmarks are close in index (VO / chain-like BAL visibility: C5 `local`), S is
// block-banded and a banded Cholesky costs n * (6 w)^2 flops instead of PCG iterations that do not converge
// on long chains (block-Jacobi PCG needs O(chain length) iterations).  One CTA walks the block columns; the
// active (w+1) x (w+1) block window lives in shared memory as a ring, the forward substitution is fused into
// the factorisation, the backward substitution streams L back from L2.  Deterministic (no atomics).
#include "ba_large.cuh"

namespace b2 {
namespace {

constexpr int kBandThreads = 256;
constexpr int kBandAhead = 3;      // block columns loaded ahead of the active window

// band block (j, k), k = 0..w, is A_{j+k, j} (6 x 6 row-major): the lower band, by block column.
// A_{j+k, j}[r][c] = S(j, j+k)[c][r] (S is stored as its upper blocks).
__global__ void large_band_fill_kernel(LargeData d) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;
  const int W1 = d.bw + 1;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  const size_t total = (size_t)d.nfree * W1 * 36;
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
    d.band[i] = v;
  }
}

__global__ void __launch_bounds__(kBandThreads) large_band_chol_kernel(LargeData d) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;
  extern __shared__ __align__(16) double bsm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int N = d.nfree, w = d.bw, W1 = w + 1;
  const int SLOT = W1 * 36 + 6;                       // one block column (w+1 blocks) + its right-hand side
  const int RS = W1 + kBandAhead;                      // ring slots: the active window + columns loaded ahead
  double* ring = bsm;
  double* xs = ring + (size_t)RS * SLOT;               // backward pass: ring of the last W1 solution blocks
  double* Lsm = xs + 6 * W1;                           // factor of the current diagonal block (36)
  __shared__ int bad;
  if (tid == 0) bad = 0;
  const double* gred = d.Sg + (size_t)d.nblk * 36;

  auto load_col = [&](int j) {                         // global band column j (+ rhs = -g_red) -> ring slot
    double* s = ring + (size_t)(j % RS) * SLOT;
    const double* g = d.band + (size_t)j * W1 * 36;
    for (int i = tid; i < W1 * 36; i += kBandThreads) s[i] = g[i];
    if (tid < 6) s[W1 * 36 + tid] = -gred[6 * j + tid];
  };
  for (int j = 0; j < min(RS - 1, N); ++j) load_col(j);
  __syncthreads();

  for (int j = 0; j < N; ++j) {
    double* cj = ring + (size_t)(j % RS) * SLOT;
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
    const int wj = min(w, N - 1 - j);                  // blocks below the diagonal in this column
    // ---- panel rows: L_{j+k,j} = A_{j+k,j} L_jj^-T, one thread per row; thread 6*wj: y_j = L_jj^-1 b_j
    if (tid < 6 * wj) {
      double* row = cj + 36 + 6 * tid;                 // row (tid % 6) of block 1 + tid / 6
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
    } else if (tid == 6 * wj) {
      double* b = cj + W1 * 36;
      double y[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        double v = b[c];
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (k < c) v = fma(-L[c][k], y[k], v);
        y[c] = v * inv[c];
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) b[c] = y[c];
#pragma unroll
      for (int r = 0; r < 6; ++r) {
#pragma unroll
        for (int c = 0; c < 6; ++c) Lsm[6 * r + c] = c <= r ? L[r][c] : 0.0;   // (block 0 of cj is still being read)
        d.binv[6 * (size_t)j + r] = inv[r];
      }
    }
    __syncthreads();
    // ---- trailing update inside the window: A_{j+a,j+b} -= L_{j+a,j} L_{j+b,j}^T (a >= b >= 1) and the
    // right-hand sides b_{j+a} -= L_{j+a,j} y_j; work items = 6x6 entries, flattened over (b, a, r, c)
    const int npair = wj * (wj + 1) / 2;
    for (int id = tid; id < npair * 36 + 6 * wj; id += kBandThreads) {
      if (id < npair * 36) {
        const int e = id % 36, pr = id / 36;
        int b = 1, rem = pr, len = wj;                 // pairs enumerated b = 1..wj, a = b..wj
        while (rem >= len) { rem -= len; --len; ++b; }
        const int a = b + rem, r = e / 6, c = e % 6;
        const double* la = cj + 36 * a + 6 * r;
        const double* lb = cj + 36 * b + 6 * c;
        double* dst = ring + (size_t)((j + b) % RS) * SLOT + 36 * (a - b) + 6 * r + c;
        double v = *dst;
#pragma unroll
        for (int m = 0; m < 6; ++m) v = fma(-la[m], lb[m], v);
        *dst = v;
      } else {
        const int q = id - npair * 36, a = 1 + q / 6, r = q % 6;
        const double* la = cj + 36 * a + 6 * r;
        const double* y = cj + W1 * 36;
        double* dst = ring + (size_t)((j + a) % RS) * SLOT + W1 * 36 + r;
        double v = *dst;
#pragma unroll
        for (int m = 0; m < 6; ++m) v = fma(-la[m], y[m], v);
        *dst = v;
      }
    }
    // ---- column j is final: L blocks back to the band array, y_j to dc; then its slot takes column j+W1
    {
      double* g = d.band + (size_t)j * W1 * 36;
      for (int i = tid; i < W1 * 36; i += kBandThreads) g[i] = i < 36 ? Lsm[i] : cj[i];
      if (tid < 6) d.dc[6 * j + tid] = cj[W1 * 36 + tid];
    }
    // the slot of column j-1 is free since the previous barrier: it takes column j-1+RS, which is first
    // touched kBandAhead steps from now (its global load latency is off the critical path)
    if (j + RS - 1 < N) load_col(j + RS - 1);
    __syncthreads();
  }
  if (bad) { if (tid == 0) ctrl->step_bad = 1; return; }

  // ---- backward substitution L^T x = y by warp 0; x blocks of the last w columns stay in shared memory.
  // Lane layout: c = lane % 6 (component), sl = lane / 6 (a-slice, 5 slices; lanes 30, 31 idle).  The entries a
  // lane needs from column j-1 (its first a-slice, the diagonal block column c, 1/diag, y) are loaded one step
  // ahead, so the L2 latency of the band array is off the serial chain.
  if (warp == 0) {
    const int c = lane % 6, sl = lane / 6;
    double nb[6], nd[6], ninv = 0.0, ny = 0.0;          // prefetched for the column about to be processed
    auto fetch = [&](int j) {
      const double* g = d.band + (size_t)j * W1 * 36;
      const int wj = min(w, N - 1 - j);
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        nb[r] = (lane < 30 && 1 + sl <= wj) ? g[36 * (1 + sl) + 6 * r + c] : 0.0;
        nd[r] = lane < 6 ? g[6 * r + c] : 0.0;          // column c of L_jj: L[r][c], r > c used
      }
      ninv = lane < 6 ? d.binv[6 * (size_t)j + lane] : 0.0;
      ny = lane < 6 ? d.dc[6 * j + lane] : 0.0;
    };
    if (N > 0) fetch(N - 1);
    for (int j = N - 1; j >= 0; --j) {
      const int wj = min(w, N - 1 - j);
      const double* g = d.band + (size_t)j * W1 * 36;
      double blk[6], dg[6];
#pragma unroll
      for (int r = 0; r < 6; ++r) { blk[r] = nb[r]; dg[r] = nd[r]; }
      const double inv_c = ninv, yc = ny;
      if (j > 0) fetch(j - 1);
      // t[c] = sum_a sum_r L_{j+a,j}[r][c] x_{j+a}[r]
      double t = 0.0;
      if (lane < 30 && 1 + sl <= wj) {
        const double* xa = xs + 6 * ((j + 1 + sl) % W1);
#pragma unroll
        for (int r = 0; r < 6; ++r) t = fma(blk[r], xa[r], t);
        for (int a = 6 + sl; a <= wj; a += 5) {         // bandwidths > 5: further slices, loaded in place
          const double* bl2 = g + 36 * a;
          const double* xa2 = xs + 6 * ((j + a) % W1);
#pragma unroll
          for (int r = 0; r < 6; ++r) t = fma(bl2[6 * r + c], xa2[r], t);
        }
      }
      t += __shfl_down_sync(0xffffffffu, t, 6) + __shfl_down_sync(0xffffffffu, t, 12) +
           __shfl_down_sync(0xffffffffu, t, 18) + __shfl_down_sync(0xffffffffu, t, 24);
      // lanes 0..5: solve L_jj^T x = y - t by back substitution, one shuffle per component
      double xc = yc - t;                               // lane c holds component c
#pragma unroll
      for (int i = 5; i >= 0; --i) {
        const double xi = __shfl_sync(0xffffffffu, xc * inv_c, i);   // x_i is final when its turn comes
        if (lane == i) xc = xi;
        else if (lane < i) xc = fma(-dg[i], xi, xc);    // L[i][c] of lane c = dg[i]
      }
      if (lane < 6) { xs[6 * (j % W1) + lane] = xc; d.dc[6 * j + lane] = xc; }
      __syncwarp();
    }
  }
  if (tid == 0) ctrl->pcg_last = 0;
}

}  // namespace

size_t large_band_smem(int bw) {
  const size_t W1 = (size_t)bw + 1;
  return ((W1 + kBandAhead) * (W1 * 36 + 6) + W1 * 6 + 36) * sizeof(double);
}

int large_band_solve(const LargeData& d, cudaStream_t s) {
  static size_t configured = 0;
  const size_t smem = large_band_smem(d.bw);
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(large_band_chol_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    configured = smem;
  }
  large_band_fill_kernel<<<296, 256, 0, s>>>(d);
  large_band_chol_kernel<<<1, kBandThreads, smem, s>>>(d);
  return (int)cudaGetLastError();
}

}  // namespace b2
