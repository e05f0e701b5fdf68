// Large path, dense reduced camera system (north_star item 4): exact blocked Cholesky of
// S dc = -g_red with the trailing updates on the FP64 tensor cores (mma.sync m8n8k4 f64, SASS DMMA).
//
//   A          (npad + 64) x npad doubles, row-major, lower triangle; rows npad.. hold the right-hand side
//              (row npad = -g_red^T), so the forward substitution y = L^-1 b falls out of the factorisation:
//              the panel's triangular solve and the trailing update treat it like any other row.
//   panel      64 columns: every CTA factors the 64 x 64 diagonal block redundantly in shared memory (8-column
//              blocks, the 8 x 8 pivot block in registers), then solves X L11^T = A21 for its 64 rows, one row per
//              thread in registers; one extra CTA produces L11^-1 for the backward pass.
//   update     C[i][j] -= sum_m A[i][k0+m] A[j][k0+m] on 128 x 128 tiles, 16 warps x (32 x 32), K chunks of 32
//              through a cp.async double buffer, C loaded straight into the DMMA accumulators.
//              Two panels (K = 128) are folded into one trailing update; the 64-column strip between them gets
//              its own narrow update.
//   look-ahead the panel chain (panels and the strips of the trailing update the next panels need) runs on a
//              high-priority stream, the bulk of every trailing update on the caller's stream underneath it.
//   backward   L^T x = y: one CTA per 64-block, resident together (cooperative launch); block k publishes x_k
//              through a release/acquire flag, block j < k folds L[k][j]^T x_k into its y_j (the L block
//              prefetched by cp.async one step ahead) and ends with x_j = L_jj^-T y_j.
#include "ba_large.cuh"

#include <cstdlib>

namespace b2 {
namespace {

constexpr int NB = 64;            // panel width
constexpr int LDL = NB + 1;       // shared-memory leading dimension of the panel matrices
constexpr int PT = 256;           // panel threads
constexpr int RPC = 64;           // rows per panel CTA (one row per thread in the triangular solve)
constexpr int TL = 128;           // update tile
constexpr int KC = 32;            // update K chunk
constexpr int LDK = KC + 4;       // == 4 mod 16: conflict-free DMMA fragment loads
constexpr int UT = 512;           // update threads (16 warps, 4 x 4)
constexpr int BT = 256;           // backward threads

__device__ __forceinline__ void cp_async16(void* dst, const void* src, int bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(src), "r"(bytes));
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// ---------------------------------------------------------------- fill
__global__ void dense_zero_kernel(const LargeCtrl* ctrl, double* A, int ld) {
  if (ctrl->stop || ctrl->step_bad) return;
  if (blockIdx.x > blockIdx.y) return;     // strictly upper 64-blocks are never read
  double2* base = reinterpret_cast<double2*>(A + (size_t)blockIdx.y * NB * ld + (size_t)blockIdx.x * NB);
  for (int q = threadIdx.x; q < NB * NB / 2; q += blockDim.x) {
    const int r = q / (NB / 2), c = q % (NB / 2);
    base[(size_t)r * (ld / 2) + c] = make_double2(0.0, 0.0);
  }
}

__global__ void dense_fill_kernel(LargeData d) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  const int ld = d.dense_ld, n = d.n;
  double* A = d.dense;
  const size_t total = (size_t)d.nblk * 36;
  for (size_t i = gid; i < total; i += gsz) {
    const int q = (int)(i / 36), e = (int)(i % 36), r = e / 6, c = e % 6;
    const int a = d.blk_row[q], b = d.blk_col[q];
    if (a == b && r > c) continue;           // diagonal block: the Schur pass maintains its upper triangle
    A[(size_t)(6 * b + c) * ld + 6 * a + r] = d.Sg[i];
  }
  const double* gred = d.Sg + (size_t)d.nblk * 36;
  for (size_t i = gid; i < (size_t)ld; i += gsz) {
    if ((int)i >= n) A[i * ld + i] = 1.0;    // identity padding up to the block size
    else A[(size_t)ld * ld + i] = -gred[i];  // right-hand side row
  }
}

// ---------------------------------------------------------------- panel
// In-place Cholesky of the 64 x 64 lower matrix L (shared memory), 8-column blocks: the 8 x 8 pivot block is
// factored redundantly in every thread's registers, the rows below solve against it one row per thread, the
// trailing update is spread over all threads.  invd = 1 / diag(L).  Returns false on a non-positive pivot.
__device__ bool potrf64(double (*L)[LDL], double* invd, int tid) {
  bool ok = true;
  for (int j0 = 0; j0 < NB; j0 += 8) {
    double D[8][8], dinv[8];
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int c = 0; c <= r; ++c) D[r][c] = L[j0 + r][j0 + c];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const double p = D[c][c];
      ok = ok && (p > 0.0);
      const double inv = rsqrt_pos(p);
      dinv[c] = inv;
      D[c][c] = p * inv;
#pragma unroll
      for (int r = c + 1; r < 8; ++r) D[r][c] *= inv;
#pragma unroll
      for (int r = c + 1; r < 8; ++r)
#pragma unroll
        for (int m = c + 1; m <= r; ++m) D[r][m] -= D[r][c] * D[m][c];
    }
    __syncthreads();                        // every thread has read the pivot block
    const int below = NB - j0 - 8;
    if (tid < below) {
      const int i = j0 + 8 + tid;
      double x[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) x[c] = L[i][j0 + c];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        x[c] *= dinv[c];
#pragma unroll
        for (int m = c + 1; m < 8; ++m) x[m] -= x[c] * D[m][c];
      }
#pragma unroll
      for (int c = 0; c < 8; ++c) L[i][j0 + c] = x[c];
    } else if (tid >= 64 && tid < 72) {
      const int r = tid - 64;
#pragma unroll
      for (int rr = 0; rr < 8; ++rr)
        if (rr == r) {
#pragma unroll
          for (int c = 0; c <= rr; ++c) L[j0 + rr][j0 + c] = D[rr][c];
          invd[j0 + rr] = dinv[rr];
        }
    }
    __syncthreads();
    // trailing update: (i, m), j0 + 8 <= m <= i < 64, 16 x 16 thread grid
    const int ti = tid >> 4, tm = tid & 15;
    for (int i = j0 + 8 + ti; i < NB; i += 16) {
      double li[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) li[c] = L[i][j0 + c];
      for (int m = j0 + 8 + tm; m <= i; m += 16) {
        double acc = L[i][m];
#pragma unroll
        for (int c = 0; c < 8; ++c) acc -= li[c] * L[m][j0 + c];
        L[i][m] = acc;
      }
    }
    __syncthreads();
  }
  return ok;
}

// blockIdx.x < nrb: rows [row_begin + RPC b, + RPC) of columns [c0, c0 + 64) := A21 L11^-T;  blockIdx.x == nrb: L11^-1.
// Every CTA factors the diagonal block itself (a latency-bound chain: repeating it costs no time and saves a
// launch boundary); RPC rows per CTA keep the number of SMs the chain takes away from the bulk update small.
__global__ void __launch_bounds__(PT) dense_panel_kernel(LargeCtrl* ctrl, double* A, int ld, int c0, int row_begin,
                                                         int row_end, int nrb, double* winv) {
  if (ctrl->stop || ctrl->step_bad) return;
  extern __shared__ __align__(16) double psm[];
  double (*L)[LDL] = reinterpret_cast<double (*)[LDL]>(psm);
  double* invd = psm + NB * LDL;
  double (*X)[LDL] = reinterpret_cast<double (*)[LDL]>(psm + NB * LDL + NB);
  const int tid = threadIdx.x;
  const bool inverse = (int)blockIdx.x == nrb;
  const size_t row0 = (size_t)row_begin + (size_t)blockIdx.x * RPC;
  const int rows = inverse ? NB : min(RPC, row_end - (int)row0);
  for (int q = tid; q < NB * NB; q += PT) {
    const int r = q >> 6, c = q & 63;
    L[r][c] = c <= r ? A[(size_t)(c0 + r) * ld + c0 + c] : 0.0;
  }
  for (int q = tid; q < rows * NB; q += PT) {
    const int r = q >> 6, c = q & 63;
    X[r][c] = inverse ? (r == c ? 1.0 : 0.0) : A[(row0 + r) * ld + c0 + c];
  }
  __syncthreads();
  const bool ok = potrf64(L, invd, tid);
  if (!ok) {                                 // uniform: every thread factored the same pivots
    if (tid == 0) ctrl->step_bad = 1;
    return;
  }
  if (tid < rows) {                          // x L^T = a, right-looking, the row in registers
    double a[NB];
#pragma unroll
    for (int c = 0; c < NB; ++c) a[c] = X[tid][c];
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      a[j] *= invd[j];
#pragma unroll
      for (int m = j + 1; m < NB; ++m) a[m] -= a[j] * L[m][j];
    }
#pragma unroll
    for (int c = 0; c < NB; ++c) X[tid][c] = a[c];
  }
  __syncthreads();
  if (inverse) {                             // X[r][m] = (L^-T)[r][m] = (L^-1)[m][r]  ->  winv = L^-1 row-major
    for (int q = tid; q < NB * NB; q += PT) {
      const int m = q >> 6, r = q & 63;
      winv[q] = X[r][m];
    }
  } else {
    for (int q = tid; q < rows * NB; q += PT) {
      const int r = q >> 6, c = q & 63;
      A[(row0 + r) * ld + c0 + c] = X[r][c];
    }
  }
}
constexpr size_t kPanelSmem = (NB * LDL + NB + RPC * LDL) * sizeof(double);

// ---------------------------------------------------------------- trailing update (DMMA)
struct UpdArgs {
  double* A;
  int ld, i_begin, i_end, j_begin, j_end, k_begin, k_len;
};

__global__ void __launch_bounds__(UT, 1) dense_update_kernel(const LargeCtrl* ctrl, UpdArgs u) {
  if (ctrl->stop || ctrl->step_bad) return;
  const int i0 = u.i_begin + TL * (int)blockIdx.y, j0 = u.j_begin + TL * (int)blockIdx.x;
  if (j0 > i0 + TL - 1) return;              // tile entirely above the diagonal
  extern __shared__ __align__(16) double usm[];
  double* Ps = usm;                          // [2][TL][LDK]
  double* Qs = usm + 2 * TL * LDK;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wi = warp >> 2, wj = warp & 3;
  const int iw0 = i0 + 32 * wi, jw0 = j0 + 32 * wj;
  const bool active = iw0 < u.i_end && jw0 < u.j_end && jw0 <= iw0 + 31;
  double* A = u.A;
  const size_t ld = (size_t)u.ld;

  auto load_chunk = [&](int ch, int st) {
    const int kb = u.k_begin + ch * KC;
#pragma unroll
    for (int x = 0; x < TL * (KC / 2) / UT; ++x) {
      const int q = tid + UT * x, row = q >> 4, cc = q & 15;
      const int gi = i0 + row, gj = j0 + row;
      const bool oki = gi < u.i_end, okj = gj < u.j_end;
      cp_async16(Ps + (size_t)(st * TL + row) * LDK + 2 * cc, A + (size_t)(oki ? gi : i0) * ld + kb + 2 * cc, oki ? 16 : 0);
      cp_async16(Qs + (size_t)(st * TL + row) * LDK + 2 * cc, A + (size_t)(okj ? gj : j0) * ld + kb + 2 * cc, okj ? 16 : 0);
    }
    cp_commit();
  };

  load_chunk(0, 0);
  double acc[4][4][2];
  if (active) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const double2 v = *reinterpret_cast<const double2*>(A + (size_t)(iw0 + 8 * a + (lane >> 2)) * ld + jw0 + 8 * b + 2 * (lane & 3));
        acc[a][b][0] = v.x;
        acc[a][b][1] = v.y;
      }
  }
  const int nch = u.k_len / KC;
  for (int ch = 0; ch < nch; ++ch) {
    if (ch + 1 < nch) { load_chunk(ch + 1, (ch + 1) & 1); cp_wait<1>(); }
    else cp_wait<0>();
    __syncthreads();
    if (active) {
      const int st = ch & 1;
      const double* pa = Ps + (size_t)(st * TL + 32 * wi + (lane >> 2)) * LDK + (lane & 3);
      const double* qb = Qs + (size_t)(st * TL + 32 * wj + (lane >> 2)) * LDK + (lane & 3);
#pragma unroll
      for (int kk = 0; kk < KC; kk += 4) {
        double a[4], b[4];
#pragma unroll
        for (int x = 0; x < 4; ++x) { a[x] = -pa[8 * x * LDK + kk]; b[x] = qb[8 * x * LDK + kk]; }
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
          for (int y = 0; y < 4; ++y) dmma884(acc[x][y][0], acc[x][y][1], a[x], b[y]);
      }
    }
    __syncthreads();
  }
  if (active) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b)
        *reinterpret_cast<double2*>(A + (size_t)(iw0 + 8 * a + (lane >> 2)) * ld + jw0 + 8 * b + 2 * (lane & 3)) =
            make_double2(acc[a][b][0], acc[a][b][1]);
  }
}
constexpr size_t kUpdateSmem = (size_t)4 * TL * LDK * sizeof(double);

// ---------------------------------------------------------------- backward substitution
__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

// out[c] = sum_r M[r][c] v[r]  (M 64 x 64 row-major in shared memory), all BT threads; result valid for tid < 64
__device__ __forceinline__ double colsum64(const double* M, const double* v, double (*part)[NB], int tid) {
  const int c = tid & 63, q = tid >> 6;
  double s = 0.0;
#pragma unroll
  for (int r = 0; r < 16; ++r) s += M[(16 * q + r) * NB + c] * v[16 * q + r];
  part[q][c] = s;
  __syncthreads();
  double t = 0.0;
  if (tid < NB) t = (part[0][c] + part[1][c]) + (part[2][c] + part[3][c]);
  return t;
}

__global__ void __launch_bounds__(BT) dense_backward_kernel(const LargeCtrl* ctrl, const double* A, int ld, int npad,
                                                            const double* winv, double* xg, int* flags, int epoch,
                                                            double* dc, int n) {
  if (ctrl->stop || ctrl->step_bad) return;
  extern __shared__ __align__(16) double bsm[];
  double* W = bsm;                           // L_jj^-1
  double* Lb = bsm + NB * NB;                // [2][64][64]
  double* yv = bsm + 3 * NB * NB;
  double* xv = yv + NB;
  double (*part)[NB] = reinterpret_cast<double (*)[NB]>(xv + NB);
  const int tid = threadIdx.x, j = blockIdx.x, nb = gridDim.x;
  auto fetch_block = [&](const double* src, size_t stride, double* dst) {
#pragma unroll
    for (int x = 0; x < NB * NB / 2 / BT; ++x) {
      const int q = tid + BT * x, r = q >> 5, cc = q & 31;
      cp_async16(dst + r * NB + 2 * cc, src + (size_t)r * stride + 2 * cc, 16);
    }
    cp_commit();
  };
  fetch_block(winv + (size_t)j * NB * NB, NB, W);
  if (nb - 1 > j) fetch_block(A + (size_t)(nb - 1) * NB * ld + (size_t)j * NB, ld, Lb);
  if (tid < NB) yv[tid] = A[(size_t)npad * ld + (size_t)j * NB + tid];
  for (int k = nb - 1; k > j; --k) {
    const int cur = (nb - 1 - k) & 1;
    if (k - 1 > j) fetch_block(A + (size_t)(k - 1) * NB * ld + (size_t)j * NB, ld, Lb + (cur ^ 1) * NB * NB);
    if (tid == 0) while (ld_acquire(flags + k) != epoch) {}
    __syncthreads();
    if (tid < NB) xv[tid] = __ldcg(xg + (size_t)k * NB + tid);
    if (k - 1 > j) cp_wait<1>(); else cp_wait<0>();
    __syncthreads();
    const double t = colsum64(Lb + cur * NB * NB, xv, part, tid);
    if (tid < NB) yv[tid] -= t;
    __syncthreads();
  }
  cp_wait<0>();
  __syncthreads();
  const double x = colsum64(W, yv, part, tid);
  if (tid < NB) {
    xg[(size_t)j * NB + tid] = x;
    if (j * NB + tid < n) dc[j * NB + tid] = x;
    __threadfence();
  }
  __syncthreads();
  if (tid == 0) st_release(flags + j, epoch);
}
constexpr size_t kBackwardSmem = (size_t)(3 * NB * NB + 2 * NB + 4 * NB) * sizeof(double);

struct DenseRt {               // launch state of one device (streams and events belong to a device)
  bool ready = false;
  cudaStream_t chain = nullptr;                   // high priority: the panel chain runs ahead of the bulk updates
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_panel[2] = {nullptr, nullptr}, ev_rest[2] = {nullptr, nullptr};
  int sms = 0;
};
DenseRt g_rts[16];             // indexed by device ordinal
int g_epoch = 0;               // publication epochs of the backward pass: monotonic across handles and devices

int dense_runtime(DenseRt** out) {
  int dev_ = 0;
  cudaGetDevice(&dev_);
  if (dev_ < 0 || dev_ >= 16) return (int)cudaErrorInvalidDevice;
  DenseRt& g_rt = g_rts[dev_];
  *out = &g_rt;
  if (g_rt.ready) return 0;
  cudaError_t e = cudaFuncSetAttribute(dense_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPanelSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(dense_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUpdateSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(dense_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBackwardSmem);
  if (e != cudaSuccess) return (int)e;
  int lo = 0, hi = 0, dev = 0;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  if ((e = cudaStreamCreateWithPriority(&g_rt.chain, cudaStreamNonBlocking, hi)) != cudaSuccess) return (int)e;   // hi = greatest priority
  cudaEvent_t* evs[] = {&g_rt.ev_fork, &g_rt.ev_join, &g_rt.ev_panel[0], &g_rt.ev_panel[1], &g_rt.ev_rest[0], &g_rt.ev_rest[1]};
  for (cudaEvent_t* ev : evs)
    if ((e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming)) != cudaSuccess) return (int)e;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&g_rt.sms, cudaDevAttrMultiProcessorCount, dev);
  g_rt.ready = true;
  return 0;
}

void launch_update(const LargeData& d, int i_begin, int j_begin, int j_end, int k_begin, int k_len, cudaStream_t st) {
  const int i_end = d.dense_ld + NB;
  if (j_end <= j_begin || i_end <= i_begin) return;
  UpdArgs u{d.dense, d.dense_ld, i_begin, i_end, j_begin, j_end, k_begin, k_len};
  const dim3 grid((j_end - j_begin + TL - 1) / TL, (i_end - i_begin + TL - 1) / TL);
  dense_update_kernel<<<grid, UT, kUpdateSmem, st>>>(d.ctrl, u);
}

void launch_panel(const LargeData& d, int c0, cudaStream_t st) {
  const int row_begin = c0 + NB, row_end = d.dense_ld + NB, nrb = (row_end - row_begin + RPC - 1) / RPC;
  dense_panel_kernel<<<nrb + 1, PT, kPanelSmem, st>>>(d.ctrl, d.dense, d.dense_ld, c0, row_begin, row_end, nrb,
                                                      d.dense_winv + (size_t)(c0 / NB) * NB * NB);
}

}  // namespace

int large_dense_npad(int n) { return (n + NB - 1) / NB * NB; }
size_t large_dense_count(int n) { const size_t np = large_dense_npad(n); return (np + NB) * np; }
int large_dense_max_n() {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms * NB;            // the backward pass keeps one CTA per 64-block resident
}

// mode (B2BA_DENSE_MODE): 0 = one update per 64-column panel; 1 = two panels per trailing update (K = 128);
// 2 (default) = 1 + look-ahead: panel chain on a high-priority stream
int large_dense_solve(const LargeData& d, cudaStream_t st) {
  DenseRt* rt = nullptr;
  int rc = dense_runtime(&rt);
  if (rc) return rc;
  DenseRt& g_rt = *rt;
  static const int mode = getenv("B2BA_DENSE_MODE") ? atoi(getenv("B2BA_DENSE_MODE")) : 2;
  const int npad = d.dense_ld, nb = npad / NB;
  dense_zero_kernel<<<dim3(nb, nb + 1), 256, 0, st>>>(d.ctrl, d.dense, npad);
  dense_fill_kernel<<<2 * d.sms, 256, 0, st>>>(d);
  if (mode == 0) {
    for (int c0 = 0; c0 < npad; c0 += NB) {
      launch_panel(d, c0, st);
      launch_update(d, c0 + NB, c0 + NB, npad, c0, NB, st);
    }
  } else {
    const bool ahead = mode >= 2;
    cudaStream_t cs = ahead ? g_rt.chain : st;  // the panel chain; the bulk updates stay on the caller's stream
    if (ahead) {
      cudaEventRecord(g_rt.ev_fork, st);
      cudaStreamWaitEvent(cs, g_rt.ev_fork, 0);
    }
    int step = 0, pending = -1;                 // pending: event index of the last bulk update
    for (int c0 = 0; c0 < npad; c0 += 2 * NB, ++step) {
      launch_panel(d, c0, cs);
      if (c0 + NB >= npad) break;
      launch_update(d, c0 + NB, c0 + NB, c0 + 2 * NB, c0, NB, cs);       // the strip the second panel needs
      launch_panel(d, c0 + NB, cs);
      const int t0 = c0 + 2 * NB;
      if (t0 >= npad) break;
      if (!ahead) {
        launch_update(d, t0, t0, npad, c0, 2 * NB, st);
        continue;
      }
      // columns [t0, t0 + 128) feed the next two panels: chain stream; the rest runs under them on the caller's
      const int t1 = t0 + 2 * NB < npad ? t0 + 2 * NB : npad;
      cudaEventRecord(g_rt.ev_panel[step & 1], cs);
      if (pending >= 0) cudaStreamWaitEvent(cs, g_rt.ev_rest[pending], 0);   // the previous bulk update wrote these columns
      pending = -1;
      launch_update(d, t0, t0, t1, c0, 2 * NB, cs);
      if (t1 < npad) {
        cudaStreamWaitEvent(st, g_rt.ev_panel[step & 1], 0);
        launch_update(d, t1, t1, npad, c0, 2 * NB, st);
        cudaEventRecord(g_rt.ev_rest[step & 1], st);
        pending = step & 1;
      }
    }
    if (ahead) {
      cudaEventRecord(g_rt.ev_join, cs);
      cudaStreamWaitEvent(st, g_rt.ev_join, 0);
    }
  }
  {
    const double* A = d.dense;
    const LargeCtrl* ctrl = d.ctrl;
    int ld = npad, np_ = npad, n = d.n, epoch = ++g_epoch;
    const double* winv = d.dense_winv;
    double* xg = d.dense_x;
    int* flags = d.dense_flags;
    double* dc = d.dc;
    void* args[] = {&ctrl, &A, &ld, &np_, &winv, &xg, &flags, &epoch, &dc, &n};
    cudaError_t e = cudaLaunchCooperativeKernel((void*)dense_backward_kernel, dim3(nb), dim3(BT), args, kBackwardSmem, st);
    if (e != cudaSuccess) return (int)e;
  }
  return (int)cudaGetLastError();
}

}  // namespace b2
