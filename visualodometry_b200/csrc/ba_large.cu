// Large path kernels -- see ba_large.cuh.
#include "ba_large.cuh"

#include <cstdlib>

#include <cooperative_groups.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <vector>

namespace cg = cooperative_groups;

namespace b2 {

namespace {

#define B2_TRY(x) do { e = (x); if (e != cudaSuccess) goto done; } while (0)

inline int bits_for(long long n) { int b = 1; while ((1ll << b) < n && b < 62) ++b; return b; }

// ------------------------------------------------------------------------------------------
// S structure (a8): upper block pairs (fi <= fj) that share a landmark, plus all diagonals.
__global__ void count_pairs(int np, const ObsRec* __restrict__ pm, const int* __restrict__ pt_ptr,
                            const int* __restrict__ fidx, int* cnt) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= np) return;
  int k = 0;
  for (int a = pt_ptr[p]; a < pt_ptr[p + 1]; ++a) k += fidx[pm[a].idx] >= 0 ? 1 : 0;
  cnt[p] = k * (k - 1) / 2;
}

__global__ void emit_pairs(int np, int nfree, const ObsRec* __restrict__ pm, const int* __restrict__ pt_ptr,
                           const int* __restrict__ fidx, const int* __restrict__ off,
                           unsigned long long* keys) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= np) return;
  int o = off[p];
  const int a1 = pt_ptr[p + 1];
  for (int a = pt_ptr[p]; a < a1; ++a) {
    const int fa = fidx[pm[a].idx];
    if (fa < 0) continue;
    for (int b = a + 1; b < a1; ++b) {
      const int fb = fidx[pm[b].idx];
      if (fb < 0) continue;
      keys[o++] = (unsigned long long)fa * nfree + fb;    // cameras ascend within a landmark
    }
  }
}

__global__ void emit_diag(int nfree, unsigned long long* keys) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f < nfree) keys[f] = (unsigned long long)f * nfree + f;
}

// sorted unique keys -> CSR (row_ptr, col); optional payload copy
__global__ void keys_to_csr(int nkeys, int nrows, int ncols, const unsigned long long* __restrict__ keys,
                            int* row_ptr, int* col) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nkeys) return;
  const int row = (int)(keys[i] / ncols);
  col[i] = (int)(keys[i] % ncols);
  const int prev = i == 0 ? -1 : (int)(keys[i - 1] / ncols);
  for (int q = prev + 1; q <= row; ++q) row_ptr[q] = i;
  if (i == nkeys - 1)
    for (int q = row + 1; q <= nrows; ++q) row_ptr[q] = nkeys;
}

// upper blocks -> entries of the full symmetric structure
__global__ void emit_full(int nfree, const int* __restrict__ blk_ptr, const int* __restrict__ blk_col,
                          unsigned long long* keys, int* vals) {
  const int fi = blockIdx.x * blockDim.x + threadIdx.x;
  if (fi >= nfree) return;
  for (int b = blk_ptr[fi]; b < blk_ptr[fi + 1]; ++b) {
    const int fj = blk_col[b];
    // entry positions: upper block b -> slot 2b-(#diagonals before) is awkward; use two arrays halves
    keys[2 * b] = (unsigned long long)fi * nfree + fj;
    vals[2 * b] = b;
    if (fi != fj) { keys[2 * b + 1] = (unsigned long long)fj * nfree + fi; vals[2 * b + 1] = b | (1 << 31); }
    else { keys[2 * b + 1] = ~0ull; vals[2 * b + 1] = -1; }   // padding, sorts last
  }
}

__global__ void pt_of_obs(int p_lo, int p_hi, const int* __restrict__ pt_ptr, int base, int* pt) {
  const int p = p_lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= p_hi) return;
  for (int a = pt_ptr[p]; a < pt_ptr[p + 1]; ++a) pt[a - base] = p;
}

__global__ void cam_keys_of_slice(int n, const ObsRec* __restrict__ pm, unsigned* keys, int* pos) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = (unsigned)pm[i].idx;
  pos[i] = i;
}

__global__ void build_cm_slice(int n, int nc, const unsigned* __restrict__ keys, const int* __restrict__ pos,
                               const ObsRec* __restrict__ pm, const int* __restrict__ pt, ObsRec* cm,
                               int* cam_ptr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  ObsRec r = pm[pos[i]];
  r.idx = pt[pos[i]];
  cm[i] = r;
  const int c = (int)keys[i], prev = i == 0 ? -1 : (int)keys[i - 1];
  for (int q = prev + 1; q <= c; ++q) cam_ptr[q] = i;
  if (i == n - 1)
    for (int q = c + 1; q <= nc; ++q) cam_ptr[q] = n;
}

__global__ void fill_int(int* p, int n, int v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

__device__ __forceinline__ int find_blk_ix(const int* __restrict__ blk_ptr, const int* __restrict__ blk_col, int fi, int fj) {
  int lo = blk_ptr[fi], hi = blk_ptr[fi + 1] - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (blk_col[mid] < fj) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// Schur pair list: every (observation a, observation b >= a) of one landmark with both cameras free
// contributes Z_a Z_b^T to the S block (f_a, f_b).
__global__ void count_all_pairs(int p_lo, int p_hi, const ObsRec* __restrict__ pm, const int* __restrict__ pt_ptr,
                                const int* __restrict__ fidx, int* cnt) {
  const int p = p_lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= p_hi) return;
  int k = 0;
  for (int a = pt_ptr[p]; a < pt_ptr[p + 1]; ++a) k += fidx[pm[a].idx] >= 0 ? 1 : 0;
  cnt[p - p_lo] = k * (k + 1) / 2;
}

__global__ void emit_all_pairs(int p_lo, int p_hi, const ObsRec* __restrict__ pm, const int* __restrict__ pt_ptr,
                               const int* __restrict__ fidx, const int* __restrict__ blk_ptr,
                               const int* __restrict__ blk_col, const int* __restrict__ off, unsigned* keys, int2* vals) {
  const int p = p_lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= p_hi) return;
  int o = off[p - p_lo];
  const int a1 = pt_ptr[p + 1];
  for (int a = pt_ptr[p]; a < a1; ++a) {
    const int fa = fidx[pm[a].idx];
    if (fa < 0) continue;
    for (int b = a; b < a1; ++b) {
      const int fb = fidx[pm[b].idx];
      if (fb < 0) continue;
      keys[o] = (unsigned)find_blk_ix(blk_ptr, blk_col, fa, fb);   // cameras ascend within a landmark
      vals[o] = make_int2(a, b);
      ++o;
    }
  }
}

__global__ void rows_of_blocks(int nfree, const int* __restrict__ blk_ptr, int* blk_row) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= nfree) return;
  for (int b = blk_ptr[f]; b < blk_ptr[f + 1]; ++b) blk_row[b] = f;
}

// sorted keys in [0, ndom) -> ptr[0..ndom]
__global__ void keys_to_ptr(int n, int ndom, const unsigned* __restrict__ keys, int* ptr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int k = (int)keys[i], prev = i == 0 ? -1 : (int)keys[i - 1];
  for (int q = prev + 1; q <= k; ++q) ptr[q] = i;
  if (i == n - 1)
    for (int q = k + 1; q <= ndom; ++q) ptr[q] = n;
}

}  // namespace

int large_build_pairs(int p_lo, int p_hi, long long no, const ObsRec* pm, const int* pt_ptr, LargeIndex& ix, cudaStream_t s) {
  cudaError_t e = cudaSuccess;
  const int npl = p_hi - p_lo, nblk = ix.nblk;
  int *cnt = nullptr, *off = nullptr;
  unsigned *k0 = nullptr, *k1 = nullptr;
  int2* v0 = nullptr;
  void* tmp = nullptr;
  size_t tb = 0;
  int last_cnt = 0, last_off = 0;
  long long total = 0;
  ix.npairs = 0;
  B2_TRY(ix.pair_ptr.alloc((size_t)nblk + 1));
  B2_TRY(ix.blk_row.alloc((size_t)nblk));
  B2_TRY(ix.pm_pt.alloc((size_t)no));
  B2_TRY(cudaMemcpyAsync(&ix.a_lo, pt_ptr + p_lo, sizeof(int), cudaMemcpyDeviceToHost, s));
  B2_TRY(cudaMemcpyAsync(&ix.a_hi, pt_ptr + p_hi, sizeof(int), cudaMemcpyDeviceToHost, s));
  B2_TRY(cudaStreamSynchronize(s));
  if (npl > 0) pt_of_obs<<<(npl + 255) / 256, 256, 0, s>>>(p_lo, p_hi, pt_ptr, 0, ix.pm_pt.p);
  if (ix.nfree > 0) rows_of_blocks<<<(ix.nfree + 255) / 256, 256, 0, s>>>(ix.nfree, ix.blk_ptr.p, ix.blk_row.p);
  if (npl > 0) {
    B2_TRY(cudaMallocAsync(&cnt, (size_t)npl * sizeof(int), s));
    B2_TRY(cudaMallocAsync(&off, (size_t)npl * sizeof(int), s));
    count_all_pairs<<<(npl + 255) / 256, 256, 0, s>>>(p_lo, p_hi, pm, pt_ptr, ix.fidx.p, cnt);
    cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt, off, npl, s);
    B2_TRY(cudaMallocAsync(&tmp, tb, s));
    B2_TRY(cub::DeviceScan::ExclusiveSum(tmp, tb, cnt, off, npl, s));
    B2_TRY(cudaMemcpyAsync(&last_cnt, cnt + npl - 1, sizeof(int), cudaMemcpyDeviceToHost, s));
    B2_TRY(cudaMemcpyAsync(&last_off, off + npl - 1, sizeof(int), cudaMemcpyDeviceToHost, s));
    B2_TRY(cudaStreamSynchronize(s));
    B2_TRY(cudaFreeAsync(tmp, s));
    tmp = nullptr;
    total = (long long)last_cnt + last_off;
  }
  if (total >= (1ll << 31)) { e = cudaErrorInvalidValue; goto done; }
  B2_TRY(ix.pair_ab.alloc((size_t)total));
  if (total == 0) {
    fill_int<<<(nblk + 256) / 256, 256, 0, s>>>(ix.pair_ptr.p, nblk + 1, 0);
    goto done;
  }
  B2_TRY(cudaMallocAsync(&k0, (size_t)total * 4, s)); B2_TRY(cudaMallocAsync(&k1, (size_t)total * 4, s));
  B2_TRY(cudaMallocAsync(&v0, (size_t)total * sizeof(int2), s));
  emit_all_pairs<<<(npl + 255) / 256, 256, 0, s>>>(p_lo, p_hi, pm, pt_ptr, ix.fidx.p, ix.blk_ptr.p, ix.blk_col.p, off, k0, v0);
  {
    unsigned long long* vin = reinterpret_cast<unsigned long long*>(v0);
    unsigned long long* vout = reinterpret_cast<unsigned long long*>(ix.pair_ab.p);
    const int kb = bits_for(nblk);
    cub::DeviceRadixSort::SortPairs(nullptr, tb, k0, k1, vin, vout, (int)total, 0, kb, s);
    B2_TRY(cudaMallocAsync(&tmp, tb, s));
    B2_TRY(cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, vin, vout, (int)total, 0, kb, s));   // stable: landmark order kept
  }
  keys_to_ptr<<<(int)((total + 255) / 256), 256, 0, s>>>((int)total, nblk, k1, ix.pair_ptr.p);
  B2_TRY(cudaGetLastError());
  ix.npairs = total;
done:
  cudaFreeAsync(cnt, s); cudaFreeAsync(off, s); cudaFreeAsync(k0, s); cudaFreeAsync(k1, s);
  cudaFreeAsync(v0, s); cudaFreeAsync(tmp, s);
  cudaStreamSynchronize(s);
  return (int)e;
}

int large_build_structure(int nc, int np, long long no, const ObsRec* pm, const int* pt_ptr,
                          const unsigned char* h_fixed, LargeIndex& ix, cudaStream_t s) {
  (void)no;
  cudaError_t e = cudaSuccess;
  std::vector<int> hf(nc);
  int nfree = 0;
  for (int c = 0; c < nc; ++c) hf[c] = h_fixed[c] ? -1 : nfree++;
  ix.nfree = nfree;
  int *cnt = nullptr, *off = nullptr, *vals = nullptr, *vals2 = nullptr, *dnum = nullptr;
  unsigned long long *keys = nullptr, *keys2 = nullptr;
  void* tmp = nullptr;
  size_t tb = 0, t2 = 0;
  int last_cnt = 0, last_off = 0, nuniq = 0;
  long long total = 0;
  const int kb = bits_for((long long)nfree * nfree + 1);
  B2_TRY(ix.fidx.alloc(nc));
  B2_TRY(cudaMemcpyAsync(ix.fidx.p, hf.data(), nc * sizeof(int), cudaMemcpyHostToDevice, s));
  B2_TRY(cudaStreamSynchronize(s));   // hf goes out of scope at return; keep it simple
  B2_TRY(cudaMallocAsync(&cnt, (np + 1) * sizeof(int), s));
  B2_TRY(cudaMallocAsync(&off, (np + 1) * sizeof(int), s));
  B2_TRY(cudaMallocAsync(&dnum, sizeof(int), s));
  if (np > 0) {
    count_pairs<<<(np + 255) / 256, 256, 0, s>>>(np, pm, pt_ptr, ix.fidx.p, cnt);
    cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt, off, np, s);
    B2_TRY(cudaMallocAsync(&tmp, tb, s));
    B2_TRY(cub::DeviceScan::ExclusiveSum(tmp, tb, cnt, off, np, s));
    B2_TRY(cudaMemcpyAsync(&last_cnt, cnt + np - 1, sizeof(int), cudaMemcpyDeviceToHost, s));
    B2_TRY(cudaMemcpyAsync(&last_off, off + np - 1, sizeof(int), cudaMemcpyDeviceToHost, s));
    B2_TRY(cudaStreamSynchronize(s));
    B2_TRY(cudaFreeAsync(tmp, s));
    tmp = nullptr;
  }
  total = (long long)last_cnt + last_off + nfree;
  if (total >= (1ll << 31)) { e = cudaErrorInvalidValue; goto done; }
  B2_TRY(cudaMallocAsync(&keys, (total + 1) * 8, s));
  B2_TRY(cudaMallocAsync(&keys2, (total + 1) * 8, s));
  if (nfree > 0) emit_diag<<<(nfree + 255) / 256, 256, 0, s>>>(nfree, keys);
  if (np > 0) emit_pairs<<<(np + 255) / 256, 256, 0, s>>>(np, nfree, pm, pt_ptr, ix.fidx.p, off, keys + nfree);
  if (total > 0) {
    cub::DeviceRadixSort::SortKeys(nullptr, tb, keys, keys2, (int)total, 0, kb, s);
    cub::DeviceSelect::Unique(nullptr, t2, keys2, keys, dnum, (int)total, s);
    if (t2 > tb) tb = t2;
    B2_TRY(cudaMallocAsync(&tmp, tb, s));
    B2_TRY(cub::DeviceRadixSort::SortKeys(tmp, tb, keys, keys2, (int)total, 0, kb, s));
    B2_TRY(cub::DeviceSelect::Unique(tmp, tb, keys2, keys, dnum, (int)total, s));
    B2_TRY(cudaMemcpyAsync(&nuniq, dnum, sizeof(int), cudaMemcpyDeviceToHost, s));
    B2_TRY(cudaStreamSynchronize(s));
    B2_TRY(cudaFreeAsync(tmp, s));
    tmp = nullptr;
  }
  ix.nblk = nuniq;
  ix.nfull = 2 * nuniq - nfree;
  B2_TRY(ix.blk_ptr.alloc(nfree + 1)); B2_TRY(ix.blk_col.alloc(nuniq));
  B2_TRY(ix.frow_ptr.alloc(nfree + 1)); B2_TRY(ix.fcol.alloc(2 * (size_t)nuniq)); B2_TRY(ix.fsrc.alloc(2 * (size_t)nuniq));
  if (nuniq == 0) {
    fill_int<<<1, 32, 0, s>>>(ix.blk_ptr.p, nfree + 1, 0);
    fill_int<<<1, 32, 0, s>>>(ix.frow_ptr.p, nfree + 1, 0);
    goto done;
  }
  keys_to_csr<<<(nuniq + 255) / 256, 256, 0, s>>>(nuniq, nfree, nfree, keys, ix.blk_ptr.p, ix.blk_col.p);
  // full symmetric structure: 2*nblk candidate entries (diagonal duplicates padded out), sorted by (row, col)
  B2_TRY(cudaMallocAsync(&vals, 2 * (size_t)nuniq * sizeof(int), s));
  B2_TRY(cudaMallocAsync(&vals2, 2 * (size_t)nuniq * sizeof(int), s));
  B2_TRY(cudaFreeAsync(keys2, s)); keys2 = nullptr;
  B2_TRY(cudaMallocAsync(&keys2, 2 * (size_t)nuniq * 8, s));
  {
    unsigned long long* fk = nullptr;
    B2_TRY(cudaMallocAsync(&fk, 2 * (size_t)nuniq * 8, s));
    emit_full<<<(nfree + 255) / 256, 256, 0, s>>>(nfree, ix.blk_ptr.p, ix.blk_col.p, fk, vals);
    cub::DeviceRadixSort::SortPairs(nullptr, tb, fk, keys2, vals, vals2, 2 * nuniq, 0, 64, s);
    e = cudaMallocAsync(&tmp, tb, s);
    if (e == cudaSuccess) e = cub::DeviceRadixSort::SortPairs(tmp, tb, fk, keys2, vals, vals2, 2 * nuniq, 0, 64, s);
    cudaFreeAsync(fk, s);
    if (e != cudaSuccess) goto done;
  }
  keys_to_csr<<<(ix.nfull + 255) / 256, 256, 0, s>>>(ix.nfull, nfree, nfree, keys2, ix.frow_ptr.p, ix.fcol.p);
  B2_TRY(cudaMemcpyAsync(ix.fsrc.p, vals2, (size_t)ix.nfull * sizeof(int), cudaMemcpyDeviceToDevice, s));
  B2_TRY(cudaGetLastError());
done:
  cudaFreeAsync(cnt, s); cudaFreeAsync(off, s); cudaFreeAsync(dnum, s); cudaFreeAsync(keys, s);
  cudaFreeAsync(keys2, s); cudaFreeAsync(vals, s); cudaFreeAsync(vals2, s); cudaFreeAsync(tmp, s);
  cudaStreamSynchronize(s);
  return (int)e;
}

int large_build_cm_slice(int nc, int p_lo, int p_hi, const ObsRec* pm, const int* pt_ptr, ObsRec* cm,
                         int* cam_ptr, cudaStream_t s) {
  cudaError_t e = cudaSuccess;
  int lo = 0, hi = 0;
  unsigned *k0 = nullptr, *k1 = nullptr;
  int *v0 = nullptr, *v1 = nullptr, *pt = nullptr;
  void* tmp = nullptr;
  size_t tb = 0;
  int n = 0;
  B2_TRY(cudaMemcpyAsync(&lo, pt_ptr + p_lo, sizeof(int), cudaMemcpyDeviceToHost, s));
  B2_TRY(cudaMemcpyAsync(&hi, pt_ptr + p_hi, sizeof(int), cudaMemcpyDeviceToHost, s));
  B2_TRY(cudaStreamSynchronize(s));
  n = hi - lo;
  if (n == 0) {
    fill_int<<<(nc + 256) / 256, 256, 0, s>>>(cam_ptr, nc + 1, 0);
    return (int)cudaGetLastError();
  }
  B2_TRY(cudaMallocAsync(&k0, n * 4, s)); B2_TRY(cudaMallocAsync(&k1, n * 4, s));
  B2_TRY(cudaMallocAsync(&v0, n * 4, s)); B2_TRY(cudaMallocAsync(&v1, n * 4, s));
  B2_TRY(cudaMallocAsync(&pt, n * 4, s));
  pt_of_obs<<<(p_hi - p_lo + 255) / 256, 256, 0, s>>>(p_lo, p_hi, pt_ptr, lo, pt);
  cam_keys_of_slice<<<(n + 255) / 256, 256, 0, s>>>(n, pm + lo, k0, v0);
  cub::DeviceRadixSort::SortPairs(nullptr, tb, k0, k1, v0, v1, n, 0, bits_for(nc), s);
  B2_TRY(cudaMallocAsync(&tmp, tb, s));
  B2_TRY(cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, v0, v1, n, 0, bits_for(nc), s));
  build_cm_slice<<<(n + 255) / 256, 256, 0, s>>>(n, nc, k1, v1, pm + lo, pt, cm, cam_ptr);
  B2_TRY(cudaGetLastError());
done:
  cudaFreeAsync(k0, s); cudaFreeAsync(k1, s); cudaFreeAsync(v0, s); cudaFreeAsync(v1, s);
  cudaFreeAsync(pt, s); cudaFreeAsync(tmp, s);
  return (int)e;
}

// nranks > 1: the block rows this rank owns -- a contiguous range [f_lo, f_hi), the ranges of all ranks balanced by
// FULL-row entries (both triangles: a rank multiplies whole rows of S) -- and their local CSR.  Host side: the
// structure is small (<= 2 * nblk ints) and is built once per problem.
int large_build_rows(int rank, int nranks, LargeIndex& ix, cudaStream_t s) {
  const int nfree = ix.nfree;
  std::vector<int> rp(nfree + 1), col(ix.nfull), src(ix.nfull);
  cudaError_t e = cudaMemcpyAsync(rp.data(), ix.frow_ptr.p, rp.size() * 4, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess && ix.nfull) e = cudaMemcpyAsync(col.data(), ix.fcol.p, (size_t)ix.nfull * 4, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess && ix.nfull) e = cudaMemcpyAsync(src.data(), ix.fsrc.p, (size_t)ix.nfull * 4, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return (int)e;
  // cut r: the first row whose prefix of entries reaches r/nranks of the total (identical on every rank)
  auto cut = [&](int r) {
    if (r <= 0) return 0;
    if (r >= nranks) return nfree;
    const long long target = (long long)ix.nfull * r / nranks;
    int f = (int)(std::lower_bound(rp.begin(), rp.end(), (int)target) - rp.begin());
    return f > nfree ? nfree : f;
  };
  ix.f_lo = cut(rank); ix.f_hi = cut(rank + 1);
  const int nrows = ix.f_hi - ix.f_lo, e0 = rp[ix.f_lo], e1 = rp[ix.f_hi];
  ix.nlocal = e1 - e0;
  std::vector<int> lrp(nrows + 1), ldiag(nrows > 0 ? nrows : 1, 0);
  for (int i = 0; i <= nrows; ++i) lrp[i] = rp[ix.f_lo + i] - e0;
  for (int i = 0; i < nrows; ++i)
    for (int q = rp[ix.f_lo + i]; q < rp[ix.f_lo + i + 1]; ++q)
      if (col[q] == ix.f_lo + i) ldiag[i] = q - e0;
  if ((e = ix.lrow_ptr.alloc(nrows + 1)) != cudaSuccess) return (int)e;
  if ((e = ix.ldiag.alloc(nrows > 0 ? nrows : 1)) != cudaSuccess) return (int)e;
  if ((e = ix.lcol.alloc(ix.nlocal > 0 ? ix.nlocal : 1)) != cudaSuccess) return (int)e;
  if ((e = ix.lsrc.alloc(ix.nlocal > 0 ? ix.nlocal : 1)) != cudaSuccess) return (int)e;
  e = cudaMemcpy(ix.lrow_ptr.p, lrp.data(), lrp.size() * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(ix.ldiag.p, ldiag.data(), ldiag.size() * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && ix.nlocal) e = cudaMemcpy(ix.lcol.p, col.data() + e0, (size_t)ix.nlocal * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && ix.nlocal) e = cudaMemcpy(ix.lsrc.p, src.data() + e0, (size_t)ix.nlocal * 4, cudaMemcpyHostToDevice);
  return (int)e;
}

}  // namespace b2

// =============================================================================================
// LM iteration kernels.  Every kernel returns immediately once ctrl->stop is set, so the host can
// enqueue a fixed number of iterations without reading anything back.
namespace b2 {
namespace {

__device__ __forceinline__ int packed21(int r, int s) { return r * 6 - r * (r - 1) / 2 + (s - r); }

__device__ __forceinline__ int find_blk(const int* __restrict__ blk_ptr, const int* __restrict__ blk_col, int fi, int fj) {
  int lo = blk_ptr[fi], hi = blk_ptr[fi + 1] - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (blk_col[mid] < fj) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// a1+a2, camera-major: one CTA per camera -> B_c (21), g_c (6); fixed-order CTA reduction.
// Camera c also owns slot c of the cost / invalid partials; thread 0 of CTA 0 sums them last.
constexpr int kCamThreads = 128;
__global__ void __launch_bounds__(kCamThreads) large_cam_kernel(LargeData d, Params P) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->have_lin) return;
  __shared__ double red[kCamThreads / 32][32];
  const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double* pose = d.poses + 12 * (size_t)c;
  const bool isfree = d.fidx[c] >= 0;
  double acc[29];
#pragma unroll
  for (int k = 0; k < 29; ++k) acc[k] = 0.0;
  for (int a = d.cam_ptr[c] + tid; a < d.cam_ptr[c + 1]; a += kCamThreads) {
    const ObsRec rec = d.cm[a];
    const double* X = d.pts + 3 * (size_t)rec.idx;
    Proj o;
    project(P, pose, X[0], X[1], X[2], rec.u, rec.v, o);
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
    if (lane == 0) red[warp][k] = v;
  }
  __syncthreads();
  if (tid < 29) {
    double v = 0.0;
    for (int w = 0; w < kCamThreads / 32; ++w) v += red[w][tid];
    if (tid < 21) d.lin_part[(size_t)c * 21 + tid] = v;
    else if (tid < 27) d.lin_part[(size_t)d.nc * 21 + (size_t)c * 6 + tid - 21] = v;
    else d.part[(tid - 27) * d.nc + c] = v;      // per-camera cost / invalid count
  }
}

// fixed-order sum of the per-camera cost partials -> lin_part tail
__global__ void large_cam_finish_kernel(LargeData d) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->have_lin) return;
  // warp 0: cost, warp 1: invalid count; lane-strided partial sums, then a fixed shuffle tree
  const int which = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double v = 0.0;
  for (int c = lane; c < d.nc; c += 32) v += d.part[which * d.nc + c];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) d.lin_part[(size_t)d.nc * 27 + which] = v;
}

// after the {B, g_c, cost} reduction: D_c, cost bookkeeping, S := blockdiag(B') (rank 0), g_red := g_c
__global__ void large_init_S_kernel(LargeData d) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  const double* B = d.lin;
  for (size_t i = gid; i < (size_t)d.nc * 6; i += gsz) d.Dc[i] = clampd(B[(i / 6) * 21 + packed21((int)(i % 6), (int)(i % 6))]);
  if (d.nranks > 1) return;                        // every entry of S is written by the owner of its row
  for (size_t i = gid; i < d.sg_count(); i += gsz) d.Sg[i] = 0.0;
  // the grid-wide zero fill must finish before the diagonal blocks are written: second kernel below
}

__global__ void large_init_diag_kernel(LargeData d) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0) {
    if (!ctrl->have_lin) {
      ctrl->cost = d.lin[(size_t)d.nc * 27];
      ctrl->cur_ninv = d.lin[(size_t)d.nc * 27 + 1];
      if (ctrl->it == 0) ctrl->initial_cost = ctrl->cost;
      ctrl->fresh_lin = 1;
    } else ctrl->fresh_lin = 0;
    ctrl->have_lin = 1;
    ctrl->step_bad = 0;
    ctrl->sums[6] = 0.0;
  }
  if (c >= d.nc || d.nranks > 1) return;           // nranks > 1: large_sblock_rows_kernel forms the owned rows, B' included
  const int f = d.fidx[c];
  if (f < 0) return;
  const double imu = 1.0 / ctrl->mu;
  const double* B = d.lin + (size_t)c * 21;
  double* S = d.Sg + (size_t)d.blk_ptr[f] * 36;      // the diagonal block is the first of its row
  for (int r = 0; r < 6; ++r)
    for (int s = 0; s < 6; ++s) {
      double v = B[packed21(min(r, s), max(r, s))];
      if (r == s) v += clampd(v) * imu;
      S[6 * r + s] = v;
    }
  double* gred = d.Sg + (size_t)d.nblk * 36;
  for (int r = 0; r < 6; ++r) gred[6 * f + r] = d.lin[(size_t)d.nc * 21 + (size_t)c * 6 + r];
}

__device__ __forceinline__ void atomic_max_pos(double* addr, double v) {   // v >= 0
  atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}

// a1-a4 in two passes without atomics (deterministic):
//  (1) large_z_kernel, point-major, one thread per landmark of this rank's range: C_p, g_p -> F (C'^-1 = F F^T),
//      per-landmark scratch for the back-substitution, and per observation  Z_k = E_k F (6 x 3)  and
//      zy_k = Z_k F^T g_p (6)  into zbuf (24 doubles per observation; zero for fixed cameras / invalid);
//  (2) large_sblock_kernel, one group of G lanes per S block: S(fa,fb) -= sum over the block's pair list of
//      Z_a Z_b^T, and for diagonal blocks g_red(f) -= sum zy_a.  The pair list is sorted by block and keeps the
//      landmark order inside a block, the lanes' partial sums are combined by a fixed shuffle tree.
__global__ void __launch_bounds__(128) large_pt_kernel(LargeData d, Params P) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const int pi = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = pi < d.np;                    // (no early exit: warp-wide max below)
  const int p = live ? pi : 0;
  const double imu = 1.0 / ctrl->mu;
  const int a0 = d.pt_ptr[p], a1 = live ? d.pt_ptr[p + 1] : a0;
  const double X0 = d.pts[3 * (size_t)p], X1 = d.pts[3 * (size_t)p + 1], X2 = d.pts[3 * (size_t)p + 2];
  double C[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
  for (int a = a0; a < a1; ++a) {
    const ObsRec rec = d.pm[a];
    const double* pose = d.poses + 12 * (size_t)rec.idx;
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
  const double D0 = clampd(C[0]), D1 = clampd(C[3]), D2 = clampd(C[5]);
  double F[6];
  {
    const double Cd[6] = {C[0] + D0 * imu, C[1], C[2], C[3] + D1 * imu, C[4], C[5] + D2 * imu};
    if (!sym3_inv_factor(Cd, F)) {
#pragma unroll
      for (int k = 0; k < 6; ++k) F[k] = 0.0;
      if (live) ctrl->step_bad = 1;
    }
  }
  if (live) {
    double2* sc = reinterpret_cast<double2*>(d.pt_scratch + (size_t)p * 12);
    sc[0] = make_double2(F[0], F[1]); sc[1] = make_double2(F[2], F[3]); sc[2] = make_double2(F[4], F[5]);
    sc[3] = make_double2(g[0], g[1]); sc[4] = make_double2(g[2], D0); sc[5] = make_double2(D1, D2);
  }
  double gm = live ? fmax(fabs(g[0]), fmax(fabs(g[1]), fabs(g[2]))) : 0.0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) gm = fmax(gm, __shfl_xor_sync(0xffffffffu, gm, o));
  if ((threadIdx.x & 31) == 0) atomic_max_pos(&ctrl->sums[6], gm);
}

// per observation of this rank's range: Z = E F (6 x 3), zy = Z F^T g_p (6) -> zbuf
__global__ void __launch_bounds__(256) large_z_kernel(LargeData d, Params P, int a_lo, int a_hi) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const int a = a_lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= a_hi) return;
  const ObsRec ra = d.pm[a];
  const int p = d.pm_pt[a];
  double2* zo = reinterpret_cast<double2*>(d.zbuf + (size_t)a * 24);
  const double* pose = d.poses + 12 * (size_t)ra.idx;
  Proj o;
  project(P, pose, d.pts[3 * (size_t)p], d.pts[3 * (size_t)p + 1], d.pts[3 * (size_t)p + 2], ra.u, ra.v, o);
  double Z[24];
  if (d.fidx[ra.idx] < 0 || !o.valid) {
#pragma unroll
    for (int k = 0; k < 24; ++k) Z[k] = 0.0;
  } else {
    const double2* sc = reinterpret_cast<const double2*>(d.pt_scratch + (size_t)p * 12);
    const double2 s0 = sc[0], s1 = sc[1], s2 = sc[2], s3 = sc[3], s4 = sc[4];
    const double F0 = s0.x, F1 = s0.y, F2 = s1.x, F3 = s1.y, F4 = s2.x, F5 = s2.y;
    const double g0 = s3.x, g1 = s3.y, g2 = s4.x;
    const double y0 = F0 * g0, y1 = fma(F1, g0, F3 * g1), y2 = fma(F2, g0, fma(F4, g1, F5 * g2));   // y = F^T g_p
    double jp[6], jc[12], E[18];
    jac_point(P, pose, o, jp);
    jac_cam(P, o, jc);
    make_E(o.w, jc, jp, E);
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const double e0 = E[3 * r], e1 = E[3 * r + 1], e2 = E[3 * r + 2];
      Z[3 * r] = e0 * F0;
      Z[3 * r + 1] = fma(e0, F1, e1 * F3);
      Z[3 * r + 2] = fma(e0, F2, fma(e1, F4, e2 * F5));
      Z[18 + r] = fma(Z[3 * r], y0, fma(Z[3 * r + 1], y1, Z[3 * r + 2] * y2));
    }
  }
#pragma unroll
  for (int k = 0; k < 12; ++k) zo[k] = make_double2(Z[2 * k], Z[2 * k + 1]);
}

// MODE 0: every block; 1: diagonal blocks only (group index = block row); 2: off-diagonal blocks only.  Dense
// visibility has diagonal pair lists of ~observations-per-camera entries next to off-diagonal lists of a handful.
template <int G, int MODE>
__global__ void __launch_bounds__(128) large_sblock_kernel(LargeData d) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const int gi = (blockIdx.x * blockDim.x + threadIdx.x) / G, gl = threadIdx.x % G;
  bool live = gi < (MODE == 1 ? d.nfree : d.nblk);   // (no early exit: the shuffles below name the full warp)
  const int gidx = !live ? 0 : MODE == 1 ? d.blk_ptr[gi] : gi;   // the diagonal block is the first of its row
  if (MODE == 2 && live && d.blk_row[gidx] == d.blk_col[gidx]) live = false;
  const int q0 = live ? d.pair_ptr[gidx] : 0, q1 = live ? d.pair_ptr[gidx + 1] : 0;
  double acc[36], ag[6];
#pragma unroll
  for (int k = 0; k < 36; ++k) acc[k] = 0.0;
#pragma unroll
  for (int k = 0; k < 6; ++k) ag[k] = 0.0;
  for (int q = q0 + gl; q < q1; q += G) {
    const int2 ab = d.pair_ab[q];
    const double2* za = reinterpret_cast<const double2*>(d.zbuf + (size_t)ab.x * 24);
    double A[18], B[18];
#pragma unroll
    for (int k = 0; k < 9; ++k) { const double2 t = za[k]; A[2 * k] = t.x; A[2 * k + 1] = t.y; }
    if (ab.x == ab.y) {
#pragma unroll
      for (int k = 0; k < 18; ++k) B[k] = A[k];
#pragma unroll
      for (int k = 0; k < 3; ++k) { const double2 t = za[9 + k]; ag[2 * k] += t.x; ag[2 * k + 1] += t.y; }
    } else {
      const double2* zb = reinterpret_cast<const double2*>(d.zbuf + (size_t)ab.y * 24);
#pragma unroll
      for (int k = 0; k < 9; ++k) { const double2 t = zb[k]; B[2 * k] = t.x; B[2 * k + 1] = t.y; }
    }
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c)
        acc[6 * r + c] = fma(A[3 * r], B[3 * c], fma(A[3 * r + 1], B[3 * c + 1], fma(A[3 * r + 2], B[3 * c + 2], acc[6 * r + c])));
  }
  // fixed shuffle tree over the G lanes of the group
#pragma unroll
  for (int m = G >> 1; m > 0; m >>= 1) {
#pragma unroll
    for (int k = 0; k < 36; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], m);
#pragma unroll
    for (int k = 0; k < 6; ++k) ag[k] += __shfl_xor_sync(0xffffffffu, ag[k], m);
  }
  if (q1 == q0) return;
  double* S = d.Sg + (size_t)gidx * 36;
#pragma unroll
  for (int k = 0; k < 36; ++k)
    if (k % G == gl) S[k] -= acc[k];
  // a diagonal block's pair list is exactly the observation list of its camera: g_red(f) -= sum zy
  const bool isdiag = d.pair_ab[q0].x == d.pair_ab[q0].y;
  if (isdiag) {
    const int f = d.blk_row[gidx];
    double* gred = d.Sg + (size_t)d.nblk * 36;
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (k % G == gl) gred[6 * f + k] -= ag[k];
  }
}

}  // namespace
}  // namespace b2

// =============================================================================================
// Peer-memory exchange (nranks > 1).  A rank publishes epoch e by storing e into slot [its rank] of every rank's flag
// array with a system-scope release AFTER its data stores (which went to the peers' arenas over NVLink); a rank that
// has seen all flags >= e may read what the others stored for epoch e.  Epochs only grow, and every rank takes part
// in every exchange, so a rank can run at most one exchange ahead of the slowest one: buffers that are written for
// epoch e + 2 are never still being read for epoch e (the q buffer, written every CG iteration, alternates by parity).
namespace b2 {
namespace {

__device__ __forceinline__ void st_release_sys(long long* p, long long v) {
  asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ long long ld_acquire_sys(const long long* p) {
  long long v;
  asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
constexpr unsigned long long kXchgTimeoutNs = 4000000000ull;   // a rank that is 4 s late is lost: give up, do not hang

// Called by ONE thread per peer (threads 0..nranks-1 of a CTA) after the CTA's / grid's data stores are complete and
// every storing thread has executed __threadfence_system().
__device__ __forceinline__ void xchg_publish(const Peer& x, int peer, long long epoch) {
  __threadfence_system();
  st_release_sys(reinterpret_cast<long long*>(x.at(peer, x.off_flags)) + x.rank, epoch);
}
// Thread `src` waits for rank `src`.  false on timeout.
__device__ __forceinline__ bool xchg_wait(const Peer& x, int src, long long epoch) {
  const long long* f = reinterpret_cast<const long long*>(x.at(x.rank, x.off_flags)) + src;
  if (ld_acquire_sys(f) >= epoch) return true;
  const unsigned long long t0 = global_ns();
  while (ld_acquire_sys(f) < epoch) {
    if (global_ns() - t0 > kXchgTimeoutNs) return false;
    __nanosleep(64);
  }
  return true;
}

// a stand-alone exchange between two ordinary kernels: publish (everything the previous kernels of this stream
// stored into the peers is complete: kernel boundaries order it), wait, bump the epoch.  One CTA, 32 threads.
__global__ void large_xchg_kernel(LargeData d) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const Peer& x = d.peer;
  const long long e = ctrl->xepoch + 1;
  bool ok = true;
  if ((int)threadIdx.x < x.nranks) {
    xchg_publish(x, threadIdx.x, e);
    ok = xchg_wait(x, threadIdx.x, e);
  }
  ok = __all_sync(0xffffffffu, ok);
  if (threadIdx.x == 0) {
    ctrl->xepoch = e;
    if (!ok) { ctrl->xerr = 1; ctrl->stop = 1; ctrl->status = B2BA_NUMERICAL_FAILURE; }
  }
}

// nranks > 1, a1-a4: the owned FULL rows of S.  One group of G lanes per local entry (f, c):
//   S(f, c) = [f == c] B'_f  -  sum over the pair list of the upper block {f, c} of  Z_a Z_b^T  (transposed for c < f)
// and, on diagonal entries,  g_red(f) = g_c(f) - sum zy_a.  MODE 1: diagonal entries only (long lists: the camera's
// observations), MODE 2: off-diagonal entries only.  Fixed summation order (pair list order, fixed shuffle tree).
template <int G, int MODE>
__global__ void __launch_bounds__(128) large_sblock_rows_kernel(LargeData d) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const int gi = (blockIdx.x * blockDim.x + threadIdx.x) / G, gl = threadIdx.x % G;
  const int nrows = d.f_hi - d.f_lo;
  bool live = gi < (MODE == 1 ? nrows : d.nlocal);      // (no early exit: the shuffles below name the full warp)
  const int e = !live ? 0 : MODE == 1 ? d.ldiag[gi] : gi;
  const int srcw = live ? d.lsrc[e] : 0;
  const int src = srcw & 0x7fffffff;
  const bool tr = srcw < 0;
  const bool isdiag = live && d.blk_row[src] == d.blk_col[src];
  if (MODE == 2 && isdiag) live = false;
  const int q0 = live ? d.pair_ptr[src] : 0, q1 = live ? d.pair_ptr[src + 1] : 0;
  double acc[36], ag[6];
#pragma unroll
  for (int k = 0; k < 36; ++k) acc[k] = 0.0;
#pragma unroll
  for (int k = 0; k < 6; ++k) ag[k] = 0.0;
  for (int q = q0 + gl; q < q1; q += G) {
    const int2 ab = d.pair_ab[q];
    const double2* za = reinterpret_cast<const double2*>(d.zbuf + (size_t)ab.x * 24);
    double A[18], B[18];
#pragma unroll
    for (int k = 0; k < 9; ++k) { const double2 t = za[k]; A[2 * k] = t.x; A[2 * k + 1] = t.y; }
    if (ab.x == ab.y) {
#pragma unroll
      for (int k = 0; k < 18; ++k) B[k] = A[k];
#pragma unroll
      for (int k = 0; k < 3; ++k) { const double2 t = za[9 + k]; ag[2 * k] += t.x; ag[2 * k + 1] += t.y; }
    } else {
      const double2* zb = reinterpret_cast<const double2*>(d.zbuf + (size_t)ab.y * 24);
#pragma unroll
      for (int k = 0; k < 9; ++k) { const double2 t = zb[k]; B[2 * k] = t.x; B[2 * k + 1] = t.y; }
    }
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c)
        acc[6 * r + c] = fma(A[3 * r], B[3 * c], fma(A[3 * r + 1], B[3 * c + 1], fma(A[3 * r + 2], B[3 * c + 2], acc[6 * r + c])));
  }
#pragma unroll
  for (int m = G >> 1; m > 0; m >>= 1) {
#pragma unroll
    for (int k = 0; k < 36; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], m);
#pragma unroll
    for (int k = 0; k < 6; ++k) ag[k] += __shfl_xor_sync(0xffffffffu, ag[k], m);
  }
  if (!live) return;
  double* S = d.Sl + (size_t)e * 36;
  if (isdiag) {
    const int f = d.blk_row[src];
    // camera of free index f: fidx is monotone, so search upwards from f
    int c = f;
    while (d.fidx[c] != f) ++c;
    const double imu = 1.0 / ctrl->mu;
    const double* B = d.lin + (size_t)c * 21;
#pragma unroll
    for (int k = 0; k < 36; ++k)
      if (k % G == gl) {
        const int r = k / 6, s2 = k % 6;
        double v = B[packed21(min(r, s2), max(r, s2))];
        if (r == s2) v += clampd(v) * imu;
        S[k] = v - acc[k];
      }
    double* gred = d.Sg + (size_t)d.nblk * 36;
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (k % G == gl) gred[6 * f + k] = d.lin[(size_t)d.nc * 21 + (size_t)c * 6 + k] - ag[k];
  } else {
#pragma unroll
    for (int k = 0; k < 36; ++k)
      if (k % G == gl) S[tr ? (k % 6) * 6 + k / 6 : k] = -acc[k];
  }
}

// nranks > 1, banded / dense solvers: the upper blocks of the owned rows and their g_red rows go into EVERY rank's
// S array (the factorisation runs replicated).  Followed by large_xchg_kernel.
__global__ void large_push_rows_kernel(LargeData d) {
  const LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const Peer& x = d.peer;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  for (size_t i = gid; i < (size_t)d.nlocal * 36; i += gsz) {
    const int e = (int)(i / 36), k = (int)(i % 36);
    const int srcw = d.lsrc[e];
    if (srcw < 0) continue;                               // lower-triangle entry: its transpose is pushed by its owner
    const double v = d.Sl[i];
    for (int r = 0; r < x.nranks; ++r) x.at(r, x.off_sg)[(size_t)srcw * 36 + k] = v;
  }
  const double* gred = d.Sg + (size_t)d.nblk * 36;
  for (size_t i = (size_t)d.f_lo * 6 + gid; i < (size_t)d.f_hi * 6; i += gsz) {
    const double v = gred[i];
    for (int r = 0; r < x.nranks; ++r)
      if (r != x.rank) x.at(r, x.off_sg)[(size_t)d.nblk * 36 + i] = v;
  }
  __threadfence_system();
}

}  // namespace
}  // namespace b2

// =============================================================================================
// a6: PCG on the block-sparse S with a block-Jacobi (6x6) preconditioner, ONE cooperative launch.
// Each CTA keeps the whole search direction p in shared memory, so an iteration needs two grid
// syncs: [p update (redundant per CTA) + q = S p on own rows + p.q]  |  [x, r, z on own rows + r.z, r.r].
// All dot products are summed in a fixed order (per-CTA partials, then one ordered pass) so the
// iteration is deterministic for a given grid size.
namespace b2 {
namespace {

constexpr int kPcgThreads = 256;
// The distributed PCG runs ONE CTA per SM (the grid size fixes the summation order on every rank).  Threads per CTA:
// at 2 ranks a CTA streams several 288-KB rows per CG iteration and more warps mean more loads in flight (512 threads:
// 57.8 -> 54.2 us per CG iteration); from 4 ranks on a CG iteration is barriers and exchanges, and the lighter CTA
// wins (256 threads: 30.1 / 25.9 us at 4 / 8 ranks against 34.2 / 30.3 us with 512).  B2BA_PCG_DIST_THREADS overrides.
inline int pcg_dist_threads(int nranks) {
  static const int forced = [] { const char* e = std::getenv("B2BA_PCG_DIST_THREADS"); return e ? std::atoi(e) : 0; }();
  if (forced == 256 || forced == 512) return forced;
  return nranks <= 2 ? 512 : 256;
}

__device__ __forceinline__ bool inv6_spd(const double* M, double* Inv) {   // M row-major, upper used
  double L[21];   // packed lower: L[i*(i+1)/2 + j]
  for (int j = 0; j < 6; ++j) {
    double dj = M[6 * j + j];
    for (int k = 0; k < j; ++k) dj -= L[j * (j + 1) / 2 + k] * L[j * (j + 1) / 2 + k];
    if (!(dj > 0.0) || !isfinite(dj)) return false;
    dj = sqrt(dj);
    L[j * (j + 1) / 2 + j] = dj;
    for (int i = j + 1; i < 6; ++i) {
      double v = M[6 * j + i];
      for (int k = 0; k < j; ++k) v -= L[i * (i + 1) / 2 + k] * L[j * (j + 1) / 2 + k];
      L[i * (i + 1) / 2 + j] = v / dj;
    }
  }
  for (int c = 0; c < 6; ++c) {
    double e[6];
    for (int i = 0; i < 6; ++i) {
      double v = i == c ? 1.0 : 0.0;
      for (int k = 0; k < i; ++k) v -= L[i * (i + 1) / 2 + k] * e[k];
      e[i] = v / L[i * (i + 1) / 2 + i];
    }
    for (int i = 5; i >= 0; --i) {
      double v = e[i];
      for (int k = i + 1; k < 6; ++k) v -= L[k * (k + 1) / 2 + i] * e[k];
      e[i] = v / L[i * (i + 1) / 2 + i];
    }
    for (int i = 0; i < 6; ++i) Inv[6 * i + c] = e[i];
  }
  return true;
}

// ordered sum of `count` partials (stride `stride`) by one warp; result broadcast through smem
__device__ __forceinline__ double ordered_sum(const double* part, int count, int stride, double* slot) {
  __syncthreads();
  if (threadIdx.x < 32) {
    double v = 0.0;
    for (int i = threadIdx.x; i < count; i += 32) v += part[(size_t)i * stride];
    v = warp_sum(v);
    if (threadIdx.x == 0) *slot = v;
  }
  __syncthreads();
  return *slot;
}

__global__ void __launch_bounds__(kPcgThreads) large_pcg_kernel(LargeData d, int max_it, double tol) {
  cg::grid_group grid = cg::this_grid();
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;       // uniform across the grid
  extern __shared__ double psh[];                  // p (n doubles) [+ qT (n doubles), see below]
  __shared__ double red[kPcgThreads / 32][4];
  __shared__ double rowred[kPcgThreads / 32][6];
  __shared__ double red2[kPcgThreads / 32][32];
  __shared__ double bcast[4];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nwarp_cta = kPcgThreads / 32;
  const int gwarp = blockIdx.x * nwarp_cta + warp, nwarps = gridDim.x * nwarp_cta;
  const int n = d.n, nfree = d.nfree;
  // Transposed half of the symmetric mat-vec.  When a second n-vector fits into shared memory (the launch then asks for
  // 2 n doubles), a CTA adds the products S(f,c)^T p_f of its rows into qT[c] there (one row at a time, every block of a
  // row has its own c: plain read-modify-write, fixed order) and publishes qT as ONE coalesced n-vector; the vectors of
  // all CTAs are then added in CTA order.  Otherwise (large n) every product goes to tbuf and is gathered back.
  unsigned dyn_smem;
  asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_smem));
  const bool use_qt = dyn_smem >= 2u * (unsigned)n * (unsigned)sizeof(double);
  double* qT = psh + n;
  const double* S = d.Sg;
  const double* gred = d.Sg + (size_t)d.nblk * 36;
  double* part = d.part;                           // [gridDim.x * 4]
  const int bl = lane / 6, rr = lane % 6;          // lanes 0..29: (block slot, row); 30, 31 idle

  // ---- setup: block-Jacobi inverse, r = b = -g_red, x = 0, z = M^-1 r, partial r.z and b.b
  double s_rz = 0.0, s_bb = 0.0;
  for (int f = blockIdx.x * kPcgThreads + tid; f < nfree; f += gridDim.x * kPcgThreads) {
    double Inv[36];
    if (!inv6_spd(S + (size_t)d.blk_ptr[f] * 36, Inv)) {
      ctrl->step_bad = 1;
      for (int k = 0; k < 36; ++k) Inv[k] = 0.0;
    }
    double rv[6];
    for (int k = 0; k < 36; ++k) d.Minv[(size_t)f * 36 + k] = Inv[k];
    for (int r_ = 0; r_ < 6; ++r_) { rv[r_] = -gred[6 * f + r_]; d.r[6 * f + r_] = rv[r_]; d.dc[6 * f + r_] = 0.0; }
    for (int r_ = 0; r_ < 6; ++r_) {
      double v = 0.0;
      for (int s_ = 0; s_ < 6; ++s_) v += Inv[6 * r_ + s_] * rv[s_];
      d.z[6 * f + r_] = v;
      s_rz += rv[r_] * v;
      s_bb += rv[r_] * rv[r_];
    }
  }
  s_rz = warp_sum(s_rz); s_bb = warp_sum(s_bb);
  if (lane == 0) { red[warp][0] = s_rz; red[warp][1] = s_bb; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0;
    for (int w = 0; w < nwarp_cta; ++w) { a += red[w][0]; b += red[w][1]; }
    part[blockIdx.x * 4] = a; part[blockIdx.x * 4 + 1] = b;
  }
  grid.sync();
  if (ctrl->step_bad) return;                      // uniform: written before the sync
  double rz = ordered_sum(part, gridDim.x, 4, &bcast[0]);
  const double bb = ordered_sum(part + 1, gridDim.x, 4, &bcast[1]);
  double rnorm2 = bb, beta = 0.0;
  int it = 0;
  grid.sync();                                     // partials are reused below
  for (int i = tid; i < n; i += kPcgThreads) psh[i] = 0.0;
  __syncthreads();
  if (bb > 0.0) {
    for (; it < max_it; ++it) {
      if (rnorm2 <= tol * tol * bb) break;
      // ---- phase A: p = z + beta p (whole vector, every CTA), q = S p on own rows, p.q
      for (int i = tid; i < n; i += kPcgThreads) psh[i] = d.z[i] + beta * psh[i];
      __syncthreads();
      // q = S p with every stored (upper) block read ONCE.  Pass 1, one CTA per block row f: its warps stride over the
      // row's upper blocks (5 block slots per warp and step): the row's own part of q_f = sum_c S(f,c) p_c, and per
      // off-diagonal block the transposed product S(f,c)^T p_f (6 doubles) into tbuf.  Pass 2 (after a grid sync):
      // q_c += the tbuf entries of column c, i.e. of the transposed entries of row c of the full index.  Rows are
      // dealt out in serpentine order (long and short rows alternate per CTA): pass 1 rows get shorter with f, pass 2
      // lists longer.
      // Odd CG iterations walk the passes backwards: what the previous iteration read last is still in L2.
      const int G_ = (int)gridDim.x, J_ = (nfree + G_ - 1) / G_;
      if (use_qt) {
        for (int i = tid; i < n; i += kPcgThreads) qT[i] = 0.0;
        __syncthreads();
      }
      for (int jj = 0; jj < J_; ++jj) {
        const int j = (it & 1) ? J_ - 1 - jj : jj;
        const int f = j * G_ + ((j & 1) ? G_ - 1 - (int)blockIdx.x : (int)blockIdx.x);
        if (f >= nfree) continue;                       // uniform per CTA
        const double* pf = psh + 6 * f;
        double v = 0.0;
        if (lane < 30) {
          for (int b = d.blk_ptr[f] + warp * 5 + bl; b < d.blk_ptr[f + 1]; b += 5 * nwarp_cta) {
            const double* M = S + (size_t)b * 36;
            const int c = d.blk_col[b];
            const double* x = psh + 6 * c;
            double t = 0.0;
#pragma unroll
            for (int s_ = 0; s_ < 6; ++s_) {
              v += M[6 * rr + s_] * x[s_];
              t += M[6 * s_ + rr] * pf[s_];
            }
            if (c != f) {
              if (use_qt) qT[6 * c + rr] += t;          // (c, rr) is this lane's alone within the row
              else d.tbuf[(size_t)b * 6 + rr] = t;
            }
          }
        }
        v += __shfl_down_sync(0xffffffffu, v, 6) + __shfl_down_sync(0xffffffffu, v, 12) +
             __shfl_down_sync(0xffffffffu, v, 18) + __shfl_down_sync(0xffffffffu, v, 24);
        if (lane < 6) rowred[warp][lane] = v;
        __syncthreads();
        if (warp == 0 && lane < 6) {
          double t = 0.0;
#pragma unroll
          for (int w = 0; w < nwarp_cta; ++w) t += rowred[w][lane];    // fixed order
          d.q[6 * f + lane] = t;
        }
        __syncthreads();
      }
      if (use_qt) {
        double* mine = d.tbuf + (size_t)blockIdx.x * n;
        for (int i = tid; i < n; i += kPcgThreads) mine[i] = qT[i];
      }
      grid.sync();
      double s_pq = 0.0;
      if (use_qt) {
        // q_i = own part + the transposed parts of all CTAs in CTA order: a CTA finishes a contiguous slice of q, 32
        // elements at a time; warp w adds the vectors of CTAs w, w + 8, ... (coalesced), warp 0 the 8 partial sums
        const int E = (n + G_ - 1) / G_, lo = (int)blockIdx.x * E, hi = min(n, lo + E);
        for (int i0 = lo; i0 < hi; i0 += 32) {
          const int i = i0 + lane;
          double a = 0.0;
          if (i < hi)
            for (int g = warp; g < G_; g += nwarp_cta) a += d.tbuf[(size_t)g * n + i];
          red2[warp][lane] = a;
          __syncthreads();
          if (warp == 0 && i < hi) {
            double t = d.q[i];
#pragma unroll
            for (int w = 0; w < nwarp_cta; ++w) t += red2[w][lane];      // fixed order
            d.q[i] = t;
            s_pq += t * psh[i];
          }
          __syncthreads();
        }
      } else {
      for (int j = 0; j * G_ < nfree; ++j) {
        const int f = nfree - 1 - (j * G_ + ((j & 1) ? G_ - 1 - (int)blockIdx.x : (int)blockIdx.x));   // long lists first
        if (f < 0) continue;
        double v = 0.0;
        if (lane < 30) {
          // transposed entries of row f = the entries with column < f: a prefix of the row (columns ascending)
          const int lo = d.frow_ptr[f], hi = d.frow_ptr[f + 1] - (d.blk_ptr[f + 1] - d.blk_ptr[f]);
          for (int b = lo + warp * 5 + bl; b < hi; b += 5 * nwarp_cta)
            v += d.tbuf[(size_t)(d.fsrc[b] & 0x7fffffff) * 6 + rr];
        }
        v += __shfl_down_sync(0xffffffffu, v, 6) + __shfl_down_sync(0xffffffffu, v, 12) +
             __shfl_down_sync(0xffffffffu, v, 18) + __shfl_down_sync(0xffffffffu, v, 24);
        if (lane < 6) rowred[warp][lane] = v;
        __syncthreads();
        if (warp == 0 && lane < 6) {
          double t = d.q[6 * f + lane];
#pragma unroll
          for (int w = 0; w < nwarp_cta; ++w) t += rowred[w][lane];    // fixed order
          d.q[6 * f + lane] = t;
          s_pq += t * psh[6 * f + lane];
        }
        __syncthreads();
      }
      }
      s_pq = warp_sum(s_pq);
      if (lane == 0) red[warp][0] = s_pq;
      __syncthreads();
      if (tid == 0) { double a = 0; for (int w = 0; w < nwarp_cta; ++w) a += red[w][0]; part[blockIdx.x * 4] = a; }
      grid.sync();
      const double pq = ordered_sum(part, gridDim.x, 4, &bcast[0]);
      if (!(pq > 0.0) || !isfinite(pq)) { if (blockIdx.x == 0 && tid == 0) ctrl->step_bad = 1; break; }
      const double alpha = rz / pq;
      // ---- phase B: x += alpha p, r -= alpha q, z = M^-1 r on own rows; r.z, r.r
      double s_a = 0.0, s_b = 0.0;
      for (int f = gwarp; f < nfree; f += nwarps) {
        double rn = 0.0;
        if (lane < 6) {
          d.dc[6 * f + lane] += alpha * psh[6 * f + lane];
          rn = d.r[6 * f + lane] - alpha * d.q[6 * f + lane];
          d.r[6 * f + lane] = rn;
        }
        double zv = 0.0;
#pragma unroll
        for (int s_ = 0; s_ < 6; ++s_) {
          const double rs = __shfl_sync(0xffffffffu, rn, s_);
          if (lane < 6) zv += d.Minv[(size_t)f * 36 + 6 * lane + s_] * rs;
        }
        if (lane < 6) { d.z[6 * f + lane] = zv; s_a += rn * zv; s_b += rn * rn; }
      }
      s_a = warp_sum(s_a); s_b = warp_sum(s_b);
      if (lane == 0) { red[warp][2] = s_a; red[warp][3] = s_b; }
      __syncthreads();
      if (tid == 0) {
        double a = 0, b = 0;
        for (int w = 0; w < nwarp_cta; ++w) { a += red[w][2]; b += red[w][3]; }
        part[blockIdx.x * 4 + 2] = a; part[blockIdx.x * 4 + 3] = b;
      }
      grid.sync();
      const double rz2 = ordered_sum(part + 2, gridDim.x, 4, &bcast[2]);
      rnorm2 = ordered_sum(part + 3, gridDim.x, 4, &bcast[3]);
      beta = rz2 / rz;
      rz = rz2;
    }
  }
  if (blockIdx.x == 0 && tid == 0) { ctrl->pcg_last = it; ctrl->pcg_iters += it; }
}

}  // namespace
}  // namespace b2

// =============================================================================================
// a6 on nranks > 1 GPUs: block-Jacobi PCG with S sharded by block row, ONE cooperative launch per rank, all ranks in
// lockstep.  Vectors (x, r, z, p) are replicated and stay bit-identical on every rank (same arithmetic on the same
// data in the same order; the grid size is the same on every rank); S exists only as the owned full rows (d.Sl).
//   set-up : a rank inverts the diagonal blocks of its rows and stores M_f^-1 and r_f = -g_red(f) into EVERY rank
//            (exchange 1); then z = M^-1 r, r.z and b.b everywhere.
//   per CG iteration: p = z + beta p (whole vector in shared memory, every CTA);  q_f = S(f,:) p for the owned rows,
//            stored into every rank's q buffer of this parity (exchange: the all-gather of q, 6 doubles per owned row);
//            p.q from the gathered q (every CTA, whole vector);  x, r, z and the partial r.z / r.r by CTA-strided rows;
//            ordered sums.  Two grid syncs and one peer exchange per iteration.
// Same stopping rule and the same operations as large_pcg_kernel / the oracle's PCG; the dot products are summed in a
// different (fixed) order, so iterates agree to rounding.
namespace b2 {
namespace {

template <int NT>
__global__ void __launch_bounds__(NT) large_pcg_dist_kernel(LargeData d, int max_it, double tol) {
  cg::grid_group grid = cg::this_grid();
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop || ctrl->step_bad) return;       // replicated state: uniform across the grid AND across the ranks
  extern __shared__ double psh[];                  // p (n doubles)
  __shared__ double red[NT / 32][4];
  __shared__ double rowred[NT / 32][6];
  __shared__ double bcast[4];
  __shared__ int xok;
  const Peer& X = d.peer;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int nwarp_cta = NT / 32;
  const int G_ = (int)gridDim.x;
  const int n = d.n, nfree = d.nfree, nrows = d.f_hi - d.f_lo;
  double* part = d.part;                           // [gridDim.x * 4]
  double* Minv = X.at(X.rank, X.off_minv);         // nfree x 37 (36 + ok marker), filled by the owners
  double* rvec = X.at(X.rank, X.off_r);            // n
  const int bl = lane / 6, rr = lane % 6;          // lanes 0..29: (block slot, row); 30, 31 idle
  long long epoch = ctrl->xepoch;

  // all CTAs wait for exchange `e` (CTA 0 publishes after a grid sync: every CTA's stores are done and fenced)
  auto exchange = [&](long long e) -> bool {
    __threadfence_system();
    grid.sync();
    if (blockIdx.x == 0 && tid < X.nranks) xchg_publish(X, tid, e);
    if (tid == 0) xok = 1;
    __syncthreads();
    if (tid < X.nranks && !xchg_wait(X, tid, e)) xok = 0;
    __syncthreads();
    return xok != 0;
  };
  auto fail_exchange = [&]() {
    if (blockIdx.x == 0 && tid == 0) { ctrl->xerr = 1; ctrl->stop = 1; ctrl->status = B2BA_NUMERICAL_FAILURE; ctrl->xepoch = epoch; }
  };

  // ---- set-up: owned rows -> M^-1, r = -g_red into every rank
  {
    const double* gred = d.Sg + (size_t)d.nblk * 36;
    for (int i = blockIdx.x * NT + tid; i < nrows; i += G_ * NT) {
      const int f = d.f_lo + i;
      double Inv[36];
      const bool ok = inv6_spd(d.Sl + (size_t)d.ldiag[i] * 36, Inv);
      for (int r_ = 0; r_ < X.nranks; ++r_) {
        double* M = X.at(r_, X.off_minv) + (size_t)f * 37;
        for (int k = 0; k < 36; ++k) M[k] = ok ? Inv[k] : 0.0;
        M[36] = ok ? 1.0 : 0.0;
        double* rv = X.at(r_, X.off_r) + 6 * (size_t)f;
        for (int k = 0; k < 6; ++k) rv[k] = -gred[6 * f + k];
      }
    }
  }
  if (!exchange(++epoch)) { fail_exchange(); return; }
  // every rank now holds M^-1 and r of every row: x = 0, z = M^-1 r, r.z, b.b (CTA-strided rows, ordered sums)
  double s_rz = 0.0, s_bb = 0.0, s_bad = 0.0;
  for (int f = blockIdx.x * NT + tid; f < nfree; f += G_ * NT) {
    const double* M = Minv + (size_t)f * 37;
    if (__ldcg(M + 36) == 0.0) s_bad = 1.0;
    double rv[6];
    for (int k = 0; k < 6; ++k) { rv[k] = __ldcg(rvec + 6 * f + k); d.r[6 * f + k] = rv[k]; d.dc[6 * f + k] = 0.0; }
    for (int r_ = 0; r_ < 6; ++r_) {
      double v = 0.0;
      for (int s_ = 0; s_ < 6; ++s_) v += __ldcg(M + 6 * r_ + s_) * rv[s_];
      d.z[6 * f + r_] = v;
      s_rz += rv[r_] * v;
      s_bb += rv[r_] * rv[r_];
    }
  }
  s_rz = warp_sum(s_rz); s_bb = warp_sum(s_bb); s_bad = warp_sum(s_bad);
  if (lane == 0) { red[warp][0] = s_rz; red[warp][1] = s_bb; red[warp][2] = s_bad; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0, c = 0;
    for (int w = 0; w < nwarp_cta; ++w) { a += red[w][0]; b += red[w][1]; c += red[w][2]; }
    part[blockIdx.x * 4] = a; part[blockIdx.x * 4 + 1] = b; part[blockIdx.x * 4 + 2] = c;
  }
  grid.sync();
  double rz = ordered_sum(part, G_, 4, &bcast[0]);
  const double bb = ordered_sum(part + 1, G_, 4, &bcast[1]);
  const double nbad = ordered_sum(part + 2, G_, 4, &bcast[2]);
  grid.sync();                                     // partials are reused below
  if (nbad > 0.0) {                                // a diagonal block of S is not SPD (on some rank): the step fails
    if (blockIdx.x == 0 && tid == 0) { ctrl->step_bad = 1; ctrl->xepoch = epoch; }
    return;
  }
  double rnorm2 = bb, beta = 0.0;
  int it = 0;
  for (int i = tid; i < n; i += NT) psh[i] = 0.0;
  __syncthreads();
  bool bad_pq = false;
  if (bb > 0.0) {
    for (; it < max_it; ++it) {
      if (rnorm2 <= tol * tol * bb) break;
      // ---- p = z + beta p (whole vector, every CTA)
      for (int i = tid; i < n; i += NT) psh[i] = d.z[i] + beta * psh[i];
      __syncthreads();
      // ---- q_f = S(f,:) p for the owned rows (one CTA per row, warps stride over the row's blocks, fixed order) -> all ranks
      const size_t qoff = X.off_qx + (size_t)((epoch + 1) & 1) * n;
      for (int i = blockIdx.x; i < nrows; i += G_) {
        const int f = d.f_lo + i;
        double v = 0.0;
        if (lane < 30) {
          for (int e = d.lrow_ptr[i] + warp * 5 + bl; e < d.lrow_ptr[i + 1]; e += 5 * nwarp_cta) {
            const double* M = d.Sl + (size_t)e * 36;
            const double* x = psh + 6 * d.lcol[e];
#pragma unroll
            for (int s_ = 0; s_ < 6; ++s_) v += M[6 * rr + s_] * x[s_];
          }
        }
        v += __shfl_down_sync(0xffffffffu, v, 6) + __shfl_down_sync(0xffffffffu, v, 12) +
             __shfl_down_sync(0xffffffffu, v, 18) + __shfl_down_sync(0xffffffffu, v, 24);
        if (lane < 6) rowred[warp][lane] = v;
        __syncthreads();
        if (warp == 0 && lane < 6) {
          double t = 0.0;
#pragma unroll
          for (int w = 0; w < nwarp_cta; ++w) t += rowred[w][lane];    // fixed order
          for (int r_ = 0; r_ < X.nranks; ++r_) X.at(r_, qoff)[6 * f + lane] = t;
        }
        __syncthreads();
      }
      if (!exchange(++epoch)) { fail_exchange(); return; }
      // ---- p.q from the gathered q: every CTA, whole vector, fixed order
      const double* q = X.at(X.rank, qoff);
      double s_pq = 0.0;
      for (int i = tid; i < n; i += NT) s_pq += psh[i] * __ldcg(q + i);
      s_pq = warp_sum(s_pq);
      if (lane == 0) red[warp][0] = s_pq;
      __syncthreads();
      if (tid == 0) { double a = 0; for (int w = 0; w < nwarp_cta; ++w) a += red[w][0]; bcast[0] = a; }
      __syncthreads();
      const double pq = bcast[0];
      if (!(pq > 0.0) || !isfinite(pq)) { bad_pq = true; break; }     // identical on every CTA and rank
      const double alpha = rz / pq;
      // ---- x += alpha p, r -= alpha q, z = M^-1 r by CTA-strided rows; partial r.z, r.r
      double s_a = 0.0, s_b = 0.0;
      for (int f = blockIdx.x * nwarp_cta + warp; f < nfree; f += G_ * nwarp_cta) {
        double rn = 0.0;
        if (lane < 6) {
          d.dc[6 * f + lane] += alpha * psh[6 * f + lane];
          rn = d.r[6 * f + lane] - alpha * __ldcg(q + 6 * f + lane);
          d.r[6 * f + lane] = rn;
        }
        double zv = 0.0;
#pragma unroll
        for (int s_ = 0; s_ < 6; ++s_) {
          const double rs = __shfl_sync(0xffffffffu, rn, s_);
          if (lane < 6) zv += __ldcg(Minv + (size_t)f * 37 + 6 * lane + s_) * rs;
        }
        if (lane < 6) { d.z[6 * f + lane] = zv; s_a += rn * zv; s_b += rn * rn; }
      }
      s_a = warp_sum(s_a); s_b = warp_sum(s_b);
      if (lane == 0) { red[warp][2] = s_a; red[warp][3] = s_b; }
      __syncthreads();
      if (tid == 0) {
        double a = 0, b = 0;
        for (int w = 0; w < nwarp_cta; ++w) { a += red[w][2]; b += red[w][3]; }
        part[blockIdx.x * 4 + 2] = a; part[blockIdx.x * 4 + 3] = b;
      }
      grid.sync();
      const double rz2 = ordered_sum(part + 2, G_, 4, &bcast[2]);
      rnorm2 = ordered_sum(part + 3, G_, 4, &bcast[3]);
      beta = rz2 / rz;
      rz = rz2;
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    if (bad_pq) ctrl->step_bad = 1;
    ctrl->pcg_last = it; ctrl->pcg_iters += it; ctrl->xepoch = epoch;
  }
}

}  // namespace
}  // namespace b2

namespace b2 {
namespace {

// trial cameras + camera part of the model decrease (one CTA)
__global__ void large_cam_update_kernel(LargeData d) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  __shared__ double red[32][4];
  __shared__ int bad;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) bad = 0;
  __syncthreads();
  double s0 = 0, s1 = 0, s2 = 0;
  if (!ctrl->step_bad) {
    for (int c = tid; c < d.nc; c += blockDim.x) {
      const int f = d.fidx[c];
      double* out = d.poses_tri + 12 * (size_t)c;
      const double* in = d.poses + 12 * (size_t)c;
      if (f < 0) { for (int k = 0; k < 12; ++k) out[k] = in[k]; continue; }
      double dl[6];
      for (int r = 0; r < 6; ++r) {
        dl[r] = d.dc[6 * f + r];
        if (!isfinite(dl[r])) bad = 1;
        s0 += dl[r] * dl[r] * d.Dc[6 * (size_t)c + r];
        s1 += dl[r] * d.lin[(size_t)d.nc * 21 + 6 * (size_t)c + r];
        s2 += dl[r] * dl[r];
      }
      retract_pose(in, dl, out);
    }
  }
  s0 = warp_sum(s0); s1 = warp_sum(s1); s2 = warp_sum(s2);
  if (lane == 0) { red[warp][0] = s0; red[warp][1] = s1; red[warp][2] = s2; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0, c = 0;
    for (int w = 0; w < (int)blockDim.x / 32; ++w) { a += red[w][0]; b += red[w][1]; c += red[w][2]; }
    ctrl->cam_sums[0] = a; ctrl->cam_sums[1] = b; ctrl->cam_sums[2] = c;
    if (bad) ctrl->step_bad = 1;
  }
}

// a7: back-substitution, trial landmark, trial cost; per-CTA partials (fixed-order sums later)
constexpr int kBsThreads = 128;
__global__ void __launch_bounds__(kBsThreads) large_backsub_kernel(LargeData d, Params P) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  __shared__ double red[kBsThreads / 32][6];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int p = d.p_lo + blockIdx.x * kBsThreads + tid;
  double s[6] = {0, 0, 0, 0, 0, 0};   // trial cost, dDd, gd, dn2, xn2, invalid
  if (!ctrl->step_bad && p < d.p_hi) {
    const int a0 = d.pt_ptr[p], a1 = d.pt_ptr[p + 1];
    const double X0 = d.pts[3 * (size_t)p], X1 = d.pts[3 * (size_t)p + 1], X2 = d.pts[3 * (size_t)p + 2];
    double v[3] = {0, 0, 0};
    for (int a = a0; a < a1; ++a) {
      const ObsRec rec = d.pm[a];
      const int f = d.fidx[rec.idx];
      if (f < 0) continue;
      const double* pose = d.poses + 12 * (size_t)rec.idx;
      Proj o;
      project(P, pose, X0, X1, X2, rec.u, rec.v, o);
      double jp[6], jc[12];
      jac_point(P, pose, o, jp);
      jac_cam(P, o, jc);
      const double* dc = d.dc + 6 * f;
      double t0 = 0.0, t1 = 0.0;
#pragma unroll
      for (int r = 0; r < 6; ++r) { t0 += jc[r] * dc[r]; t1 += jc[6 + r] * dc[r]; }
#pragma unroll
      for (int j = 0; j < 3; ++j) v[j] += o.w * (jp[j] * t0 + jp[3 + j] * t1);
    }
    const double* sc = d.pt_scratch + (size_t)p * 12;
    const double F0 = sc[0], F1 = sc[1], F2 = sc[2], F3 = sc[3], F4 = sc[4], F5 = sc[5];
    const double g0 = sc[6], g1 = sc[7], g2 = sc[8];
    const double t0 = g0 + v[0], t1 = g1 + v[1], t2 = g2 + v[2];
    const double u0 = F0 * t0, u1 = F1 * t0 + F3 * t1, u2 = F2 * t0 + F4 * t1 + F5 * t2;
    const double d0 = -(F0 * u0 + F1 * u1 + F2 * u2), d1 = -(F3 * u1 + F4 * u2), d2 = -(F5 * u2);
    const double Xn0 = X0 + d0, Xn1 = X1 + d1, Xn2 = X2 + d2;
    if (d.nranks > 1) {      // the trial landmark goes into EVERY rank's copy (landmarks are replicated state)
      for (int r_ = 0; r_ < d.nranks; ++r_) {
        double* pt = d.peer.at(r_, d.peer.off_pts_tri) + 3 * (size_t)p;
        pt[0] = Xn0; pt[1] = Xn1; pt[2] = Xn2;
      }
    } else {
      d.pts_tri[3 * (size_t)p] = Xn0; d.pts_tri[3 * (size_t)p + 1] = Xn1; d.pts_tri[3 * (size_t)p + 2] = Xn2;
    }
    s[1] = d0 * d0 * sc[9] + d1 * d1 * sc[10] + d2 * d2 * sc[11];
    s[2] = d0 * g0 + d1 * g1 + d2 * g2;
    s[3] = d0 * d0 + d1 * d1 + d2 * d2;
    s[4] = X0 * X0 + X1 * X1 + X2 * X2;
    for (int a = a0; a < a1; ++a) {
      const ObsRec rec = d.pm[a];
      Proj o;
      project(P, d.poses_tri + 12 * (size_t)rec.idx, Xn0, Xn1, Xn2, rec.u, rec.v, o);
      if (o.valid) s[0] += 0.5 * o.rho; else s[5] += 1.0;
    }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const double v = warp_sum(s[k]);
    if (lane == 0) red[warp][k] = v;
  }
  __syncthreads();
  if (tid < 6) {
    double v = 0.0;
    for (int w = 0; w < kBsThreads / 32; ++w) v += red[w][tid];
    d.part[(size_t)blockIdx.x * 8 + tid] = v;
  }
  if (d.nranks > 1) __threadfence_system();     // peer stores of this CTA are ordered before the exchange kernel's flag
}

// ordered sum of the back-substitution partials -> sums_part[0..5] (+ local gmax in [6])
__global__ void large_reduce_kernel(LargeData d, int nparts) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const int k = threadIdx.x / 32, lane = threadIdx.x & 31;   // 6 warps, one per sum
  if (k < 6) {
    double v = 0.0;
    for (int i = lane; i < nparts; i += 32) v += d.part[(size_t)i * 8 + k];
    v = warp_sum(v);
    if (lane == 0) d.sums_part[k] = v;
  }
  if (threadIdx.x == 0) { d.sums_part[6] = ctrl->sums[6]; d.sums_part[7] = ctrl->step_bad ? 1.0 : 0.0; }
}

// nranks > 1: the 8 partial sums of every rank's landmark range -> every rank; summed in RANK ORDER (identical
// everywhere), gmax / step-bad by max.  One CTA of 32 threads; the trial landmarks stored by large_backsub_kernel are
// covered by the same flag (kernel boundary + system fence order them before it).
__global__ void large_xchg_sums_kernel(LargeData d) {
  LargeCtrl* ctrl = d.ctrl;
  if (ctrl->stop) return;
  const Peer& x = d.peer;
  const int t = threadIdx.x;
  const long long e = ctrl->xepoch + 1;
  if (t < 8)
    for (int r = 0; r < x.nranks; ++r) x.at(r, x.off_sums)[x.rank * 8 + t] = d.sums_part[t];
  __threadfence_system();
  __syncwarp();
  bool ok = true;
  if (t < x.nranks) {
    xchg_publish(x, t, e);
    ok = xchg_wait(x, t, e);
  }
  ok = __all_sync(0xffffffffu, ok);
  if (t < 8 && ok) {
    const double* sl = x.at(x.rank, x.off_sums);
    double v = 0.0;
    for (int r = 0; r < x.nranks; ++r) {
      const double a = __ldcg(sl + r * 8 + t);
      v = (t >= 6) ? fmax(v, a) : v + a;
    }
    d.sums_part[8 + t] = v;
  }
  if (t == 0) {
    ctrl->xepoch = e;
    if (!ok) { ctrl->xerr = 1; ctrl->stop = 1; ctrl->status = B2BA_NUMERICAL_FAILURE; }
  }
}

// LM control (Spec a0), one thread; identical on every rank (replicated inputs, rank-ordered sums)
__global__ void large_control_kernel(LargeData d, Params P, const double* sums) {
  LargeCtrl* c = d.ctrl;
  if (c->stop || threadIdx.x != 0) return;
  const bool step_ok = sums[7] == 0.0 && !c->step_bad;
  const double tc = sums[0];
  const double dDd = sums[1] + c->cam_sums[0], gd = sums[2] + c->cam_sums[1], dn2 = sums[3] + c->cam_sums[2];
  const double xn2 = sums[4];
  double gm = sums[6];
  if (P.gtol > 0.0)
    for (int i = 0; i < d.nc * 6; ++i)
      if (d.fidx[i / 6] >= 0) gm = fmax(gm, fabs(d.lin[(size_t)d.nc * 21 + i]));
  const double mu = c->mu;
  const double model = 0.5 * (dDd / mu - gd);
  int accept = 0, stop = 0;
  double rho = 0.0;
  if (step_ok && model > 0.0) {
    rho = (c->cost - tc) / model;
    // a step that pushes MORE observations behind the cheirality gate (they cost 0) is rejected like a failed one
    accept = (isfinite(tc) && rho > kMinRatio && sums[5] <= c->cur_ninv) ? 1 : 0;
  }
  if (P.gtol > 0.0 && c->fresh_lin && gm <= P.gtol) { c->status = B2BA_CONVERGED; c->stop = 1; c->accept = 0; return; }
  if (accept) {
    if (P.ftol > 0.0 && fabs(c->cost - tc) <= P.ftol * c->cost) stop = B2BA_CONVERGED;
    if (P.ptol > 0.0 && sqrt(dn2) <= P.ptol * (sqrt(xn2) + P.ptol)) stop = B2BA_CONVERGED;
    const double t = 2.0 * rho - 1.0;
    c->mu = fmin(kMuMax, mu / fmax(1.0 / 3.0, 1.0 - t * t * t));
    c->nu = 2.0;
    c->cost = tc;
    c->cur_ninv = sums[5];
    c->have_lin = 0;
    c->n_acc++;
  } else {
    c->mu = mu / c->nu;
    c->nu *= 2.0;
    c->n_rej++;
  }
  c->accept = accept;
  if (d.hist_cost != nullptr && c->it < 64) {
    d.hist_cost[c->it] = c->cost; d.hist_mu[c->it] = mu; d.hist_accept[c->it] = accept;
  }
  c->it++;
  if (!stop && !(c->mu > kMuMin)) stop = B2BA_NUMERICAL_FAILURE;
  if (!stop && c->it >= P.max_iters) { stop = B2BA_MAX_ITERS; }
  if (stop) { c->status = stop; c->stop = 1; }
}

// accepted step: trial state becomes the current state (every rank keeps ALL landmarks current
// only inside its own range; the final gather happens on the host side of b2ba_get)
__global__ void large_commit_kernel(LargeData d) {
  const LargeCtrl* c = d.ctrl;
  if (!c->accept) return;
  const size_t gid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, gsz = (size_t)gridDim.x * blockDim.x;
  for (size_t i = gid; i < (size_t)d.nc * 12; i += gsz) d.poses[i] = d.poses_tri[i];
  for (size_t i = gid; i < (size_t)d.np * 3; i += gsz) d.pts[i] = d.pts_tri[i];    // landmarks are replicated state
}

__global__ void large_clear_accept_kernel(LargeData d) { d.ctrl->accept = 0; }

__global__ void large_init_ctrl_kernel(LargeData d, Params P) {
  LargeCtrl* c = d.ctrl;
  c->xerr = 0;              // (xepoch is NOT reset: peer flags only grow over the life of the communicator)
  c->mu = P.init_mu; c->nu = 2.0; c->cost = 0.0; c->initial_cost = 0.0;
  for (int k = 0; k < 8; ++k) c->sums[k] = 0.0;
  for (int k = 0; k < 4; ++k) c->cam_sums[k] = 0.0;
  c->cur_ninv = 0.0;
  c->it = 0; c->n_acc = 0; c->n_rej = 0; c->status = B2BA_MAX_ITERS; c->stop = 0; c->step_bad = 0;
  c->have_lin = 0; c->accept = 0; c->pcg_iters = 0; c->pcg_last = 0; c->fresh_lin = 0; c->pad = 0;
}

__global__ void large_finish_kernel(LargeData d) {
  const LargeCtrl* c = d.ctrl;
  b2ba_summary s;
  s.status = c->status; s.iterations = c->it; s.accepted = c->n_acc; s.rejected = c->n_rej;
  s.n_invalid = (int)c->cur_ninv; s.pcg_iters = c->pcg_iters;
  s.initial_cost = c->initial_cost; s.final_cost = c->cost; s.final_mu = c->mu;
  *d.summary = s;
}

}  // namespace

int large_pcg_grid(int n, int* blocks, size_t* smem) {
  // p, and (when it fits: n <= 12800) the CTA's transposed-product vector qT: see large_pcg_kernel.  B2BA_PCG_QT=0
  // (developer knob) keeps the tbuf gather.
  static const bool qt_off = [] { const char* e = std::getenv("B2BA_PCG_QT"); return e && std::atoi(e) == 0; }();
  *smem = (size_t)n * sizeof(double);
  if (!qt_off && 2 * *smem <= 200 * 1024) *smem *= 2;
  cudaError_t e = cudaFuncSetAttribute(large_pcg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem);
  if (e != cudaSuccess) return (int)e;
  int per_sm = 0, dev = 0, sms = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, large_pcg_kernel, kPcgThreads, *smem);
  if (e != cudaSuccess) return (int)e;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (per_sm < 1) return (int)cudaErrorInvalidConfiguration;
  static const int cap = [] { const char* e = std::getenv("B2BA_PCG_CTAS_PER_SM"); return e ? std::atoi(e) : 2; }();
  *blocks = sms * (per_sm > cap ? (cap < 1 ? 1 : cap) : per_sm);
  return 0;
}

int large_init_ctrl(const LargeData& d, const Params& P, cudaStream_t s) {
  large_init_ctrl_kernel<<<1, 1, 0, s>>>(d, P);
  return (int)cudaGetLastError();
}

int large_pcg_dist_grid(int n, int nranks, int* blocks, size_t* smem) {
  *smem = (size_t)n * sizeof(double);
  const int nt = pcg_dist_threads(nranks);
  const void* fn = nt == 512 ? (const void*)large_pcg_dist_kernel<512> : (const void*)large_pcg_dist_kernel<256>;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem);
  if (e != cudaSuccess) return (int)e;
  int per_sm = 0, dev = 0, sms = 0;
  e = nt == 512 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, large_pcg_dist_kernel<512>, 512, *smem)
                : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, large_pcg_dist_kernel<256>, 256, *smem);
  if (e != cudaSuccess) return (int)e;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (per_sm < 1) return (int)cudaErrorInvalidConfiguration;
  *blocks = sms;            // one CTA per SM: the grid size fixes the summation order, it must be the same on every rank
  return 0;
}

int large_iteration(const LargeData& d, const Params& P, int pcg_max_iters, double pcg_tol, cudaStream_t s, cudaEvent_t* ev) {
  int rc;
  const bool mr = d.nranks > 1;
  int phase = 0;
  // phase boundary: NVTX range of the host enqueue + (profiling mode) an event on the stream
  auto mark = [&](const char* next) {
    if (phase > 0) nvtxRangePop();
    if (ev) cudaEventRecord(ev[phase], s);
    if (next) nvtxRangePushA(next);
    ++phase;
  };
  mark("b2ba:linearise");
  large_clear_accept_kernel<<<1, 1, 0, s>>>(d);
  if (d.nc > 0) large_cam_kernel<<<d.nc, kCamThreads, 0, s>>>(d, P);
  large_cam_finish_kernel<<<1, 64, 0, s>>>(d);
  large_init_S_kernel<<<2 * d.sms, 256, 0, s>>>(d);
  large_init_diag_kernel<<<(d.nc + 127) / 128 + 1, 128, 0, s>>>(d);
  if (d.np > 0) {
    large_pt_kernel<<<(d.np + 127) / 128, 128, 0, s>>>(d, P);
    if (d.a_hi > d.a_lo) large_z_kernel<<<(d.a_hi - d.a_lo + 255) / 256, 256, 0, s>>>(d, P, d.a_lo, d.a_hi);
  }
  mark("b2ba:schur");
  if (mr) {
    // S sharded by block row: the owned full rows (diagonal entries carry B' and g_red), nothing to reduce
    const int nrows = d.f_hi - d.f_lo;
    if (nrows > 0) large_sblock_rows_kernel<32, 1><<<(int)(((size_t)nrows * 32 + 127) / 128), 128, 0, s>>>(d);
    if (d.nlocal > 0) {
      if (d.pair_lanes == 1) large_sblock_rows_kernel<1, 2><<<(d.nlocal + 127) / 128, 128, 0, s>>>(d);
      else if (d.pair_lanes == 8) large_sblock_rows_kernel<8, 2><<<(int)(((size_t)d.nlocal * 8 + 127) / 128), 128, 0, s>>>(d);
      else large_sblock_rows_kernel<32, 2><<<(int)(((size_t)d.nlocal * 32 + 127) / 128), 128, 0, s>>>(d);
    }
  } else if (d.np > 0) {
    // lanes per S block: dense visibility gives many blocks with a handful of pairs each
    if (d.pair_lanes == 32) {
      large_sblock_kernel<32, 0><<<(int)(((size_t)d.nblk * 32 + 127) / 128), 128, 0, s>>>(d);
    } else {
      large_sblock_kernel<32, 1><<<(int)(((size_t)d.nfree * 32 + 127) / 128), 128, 0, s>>>(d);
      if (d.pair_lanes == 1) large_sblock_kernel<1, 2><<<(d.nblk + 127) / 128, 128, 0, s>>>(d);
      else large_sblock_kernel<8, 2><<<(int)(((size_t)d.nblk * 8 + 127) / 128), 128, 0, s>>>(d);
    }
  }
  mark("b2ba:exchange");
  if (mr && (d.dense || d.bw >= 0)) {
    // factorisations run replicated: the owned rows go into every rank's S (small when S is banded)
    large_push_rows_kernel<<<2 * d.sms, 256, 0, s>>>(d);
    large_xchg_kernel<<<1, 32, 0, s>>>(d);
  }
  mark("b2ba:solve");
  if (d.dense) {
    if ((rc = large_dense_solve(d, s)) != 0) { nvtxRangePop(); return rc; }
  } else if (d.bw >= 0) {
    if ((rc = large_band_solve(d, s)) != 0) { nvtxRangePop(); return rc; }
  } else {
    int blocks = 0;
    size_t smem = 0;
    if ((rc = mr ? large_pcg_dist_grid(d.n, d.nranks, &blocks, &smem) : large_pcg_grid(d.n, &blocks, &smem)) != 0) { nvtxRangePop(); return rc; }
    LargeData dd = d;
    int mi = pcg_max_iters;
    double tol = pcg_tol;
    void* args[] = {&dd, &mi, &tol};
    const int nt = mr ? pcg_dist_threads(d.nranks) : kPcgThreads;
    void* fn = !mr ? (void*)large_pcg_kernel : nt == 512 ? (void*)large_pcg_dist_kernel<512> : (void*)large_pcg_dist_kernel<256>;
    cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(blocks), dim3(nt), args, smem, s);
    if (e != cudaSuccess) { nvtxRangePop(); return (int)e; }
  }
  mark("b2ba:backsub");
  large_cam_update_kernel<<<1, 1024, 0, s>>>(d);
  const int npl = d.p_hi - d.p_lo;
  const int nparts = (npl + kBsThreads - 1) / kBsThreads;
  if (nparts > 0) large_backsub_kernel<<<nparts, kBsThreads, 0, s>>>(d, P);
  mark("b2ba:control");
  large_reduce_kernel<<<1, 192, 0, s>>>(d, nparts);
  const double* sums = d.sums_part;
  if (mr) {
    large_xchg_sums_kernel<<<1, 32, 0, s>>>(d);
    sums = d.sums_part + 8;
  }
  large_control_kernel<<<1, 32, 0, s>>>(d, P, sums);
  large_commit_kernel<<<d.sms, 256, 0, s>>>(d);
  mark(nullptr);
  return (int)cudaGetLastError();
}

int large_finish(const LargeData& d, cudaStream_t s) {
  large_finish_kernel<<<1, 1, 0, s>>>(d);
  return (int)cudaGetLastError();
}

}  // namespace b2
