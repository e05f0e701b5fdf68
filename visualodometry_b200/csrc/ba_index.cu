// a8 index builder (see ba_index.cuh).  Two stable LSD radix sorts (CUB, part of the CUDA
// toolkit) + three small kernels.  Runs once per problem, not in the LM loop.
#include "ba_index.cuh"

#include <cub/device/device_radix_sort.cuh>

#include <cstdlib>
#include <mutex>

namespace b2 {

namespace {


// pos_of (or null): global landmark -> global schedule position; the point-major stream is then laid out in
// schedule order (the canonical order is by landmark id)
__global__ void make_pm_keys(IndexIn in, const int* __restrict__ pos_of, unsigned long long* keys, int* vals, int* flags) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= in.no) return;
  const int b = find_window(in.obs_off, in.nb, i);
  const int nc = in.cam_off[b + 1] - in.cam_off[b], np = in.pt_off[b + 1] - in.pt_off[b];
  int c = in.obs_cam[i], p = in.obs_pt[i];
  if (c < 0 || c >= nc || p < 0 || p >= np) { flags[0] = 1; c = 0; p = 0; }
  const int g = in.pt_off[b] + p;
  keys[i] = ((unsigned long long)(unsigned)(pos_of ? pos_of[g] : g) << 32) | (unsigned)c;
  vals[i] = (int)i;
}

// observations per landmark (integer atomics: the counts are exact whatever the order)
__global__ void count_obs(IndexIn in, int* cnt) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= in.no) return;
  const int b = find_window(in.obs_off, in.nb, i);
  const int np = in.pt_off[b + 1] - in.pt_off[b];
  const int p = in.obs_pt[i];
  if (p >= 0 && p < np) atomicAdd(&cnt[in.pt_off[b] + p], 1);
}

__global__ void invert_sched(IndexIn in, const int* __restrict__ sched, int* pos_of) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;   // global schedule position
  if (g >= in.np_tot) return;
  int lo = 0, hi = in.nb;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (in.pt_off[mid] <= g) lo = mid; else hi = mid; }
  pos_of[in.pt_off[lo] + sched[g]] = g;
}

// sorted (key, original index) -> point-major records, pt_ptr boundaries, camera keys for pass 2
__global__ void build_pm(IndexIn in, const unsigned long long* __restrict__ keys,
                         const int* __restrict__ vals, ObsRec* pm, int* pt_ptr, unsigned* cam_keys,
                         int* pos, int* flags) {
  const long long a = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (a >= in.no) return;
  const unsigned long long k = keys[a];
  const int gp = (int)(k >> 32), c = (int)(k & 0xffffffffu), i = vals[a];
  const float2 uv = in.obs_uv[i];
  ObsRec r; r.u = uv.x; r.v = uv.y; r.idx = c; r.orig = i;
  pm[a] = r;
  const int b = find_window(in.obs_off, in.nb, i);
  cam_keys[a] = (unsigned)(in.cam_off[b] + c);
  pos[a] = (int)a;
  const int prev = a == 0 ? -1 : (int)(keys[a - 1] >> 32);
  if (a > 0 && keys[a - 1] == k) flags[1] = 1;
  for (int q = prev + 1; q <= gp; ++q) pt_ptr[q] = (int)a;
  if (a == in.no - 1)
    for (int q = gp + 1; q <= in.np_tot; ++q) pt_ptr[q] = (int)in.no;
}

__global__ void build_cm(IndexIn in, const unsigned* __restrict__ cam_keys, const int* __restrict__ pos,
                         const ObsRec* __restrict__ pm, const unsigned long long* __restrict__ pm_keys,
                         const int* __restrict__ sched, ObsRec* cm, int* cam_ptr) {
  const long long a = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (a >= in.no) return;
  const int gc = (int)cam_keys[a];
  const int src = pos[a];
  ObsRec r = pm[src];
  const int gp = (int)(pm_keys[src] >> 32);
  const int b = find_window(in.obs_off, in.nb, r.orig);
  r.idx = gp - in.pt_off[b];      // window-local landmark slot (schedule position when the stream is schedule-ordered)
  cm[a] = r;
  const int prev = a == 0 ? -1 : (int)cam_keys[a - 1];
  for (int q = prev + 1; q <= gc; ++q) cam_ptr[q] = (int)a;
  if (a == in.no - 1)
    for (int q = gc + 1; q <= in.nc_tot; ++q) cam_ptr[q] = (int)in.no;
}

__device__ __forceinline__ int find_window_pt(const int* __restrict__ pt_off, int nb, int g) {
  int lo = 0, hi = nb;   // pt_off[lo] <= g < pt_off[hi]; empty windows are skipped by taking the LAST match
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (pt_off[mid] <= g) lo = mid; else hi = mid;
  }
  return lo;
}

// landmark schedule keys: (window, 31 - min(observation count, 31)); value = window-local landmark id
__global__ void make_sched_keys(IndexIn in, const int* __restrict__ cnt, unsigned* keys, int* vals) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= in.np_tot) return;
  const int b = find_window_pt(in.pt_off, in.nb, g);
  const int k = cnt[g];
  keys[g] = ((unsigned)b << 5) | (unsigned)(31 - min(k, 31));
  vals[g] = g - in.pt_off[b];
}

__global__ void fill_ptr(int* p, int n, int v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

inline int bits_for(long long n) { int b = 1; while ((1ll << b) < n && b < 32) ++b; return b; }

// landmark schedule (out.sched) and its inverse pos_of (np_tot ints, caller-owned): counts -> stable sort
cudaError_t sort_sched(const IndexIn& in, const IndexOut& out, int* pos_of, cudaStream_t s) {
  const int np = in.np_tot;
  unsigned *k0 = nullptr, *k1 = nullptr;
  int *v0 = nullptr, *cnt = nullptr;
  void* t = nullptr;
  size_t need = 0;
  const int bits = 5 + bits_for(in.nb);
  cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, need, k0, k1, v0, out.sched, np, 0, bits, s);
  if (e != cudaSuccess) return e;
  if ((e = cudaMallocAsync(&k0, (size_t)np * 4, s)) != cudaSuccess) goto done;
  if ((e = cudaMallocAsync(&k1, (size_t)np * 4, s)) != cudaSuccess) goto done;
  if ((e = cudaMallocAsync(&v0, (size_t)np * 4, s)) != cudaSuccess) goto done;
  if ((e = cudaMallocAsync(&cnt, (size_t)np * 4, s)) != cudaSuccess) goto done;
  if ((e = cudaMallocAsync(&t, need, s)) != cudaSuccess) goto done;
  if ((e = cudaMemsetAsync(cnt, 0, (size_t)np * 4, s)) != cudaSuccess) goto done;
  if (in.no > 0) count_obs<<<(int)((in.no + 255) / 256), 256, 0, s>>>(in, cnt);
  make_sched_keys<<<(np + 255) / 256, 256, 0, s>>>(in, cnt, k0, v0);
  e = cub::DeviceRadixSort::SortPairs(t, need, k0, k1, v0, out.sched, np, 0, bits, s);
  if (e == cudaSuccess && pos_of) invert_sched<<<(np + 255) / 256, 256, 0, s>>>(in, out.sched, pos_of);
done:
  cudaFreeAsync(k0, s); cudaFreeAsync(k1, s); cudaFreeAsync(v0, s); cudaFreeAsync(cnt, s); cudaFreeAsync(t, s);
  return e;
}

}  // namespace

namespace {
// rows of 3 doubles between landmark order and schedule order: to_sched ? dst[pos] = src[landmark] : dst[landmark] = src[pos]
__global__ void permute_rows3(int nb, int np_tot, const int* __restrict__ pt_off, const int* __restrict__ sched,
                              const double* __restrict__ src, double* dst, int to_sched) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= np_tot) return;
  const int b = find_window_pt(pt_off, nb, g);
  const size_t l = (size_t)pt_off[b] + sched[g];
  const size_t from = to_sched ? l : (size_t)g, to = to_sched ? (size_t)g : l;
  dst[3 * to] = src[3 * from]; dst[3 * to + 1] = src[3 * from + 1]; dst[3 * to + 2] = src[3 * from + 2];
}
}  // namespace

int permute_points(int nb, int np_tot, const int* pt_off, const int* sched, const double* src, double* dst,
                   bool to_sched, cudaStream_t s) {
  if (np_tot == 0) return 0;
  permute_rows3<<<(np_tot + 255) / 256, 256, 0, s>>>(nb, np_tot, pt_off, sched, src, dst, to_sched ? 1 : 0);
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// Window path: the whole index of ONE window built by ONE CTA in shared memory -- no global sort.  A window's landmark
// has every camera at most once, so the point-major order (slot, camera, original index) and the camera-major order
// (camera, slot) follow from one 32-bit camera mask per landmark: position inside a landmark's run = number of lower
// cameras that see it (popc), position inside a camera's run = number of lower slots it sees (per-camera bit vectors
// over the slots + prefix counts per 32 slots).  The landmark schedule (decreasing observation count, stable) is a
// stable counting sort over 32 buckets.  Output is bit-identical to the global radix-sort builder below (and to the
// oracle's std::stable_sort): tests/test_gpu_parity.py::test_index_bit_exact, ::test_landmark_schedule_is_stable_count_sort.
// HBM traffic: the observation arrays twice in (8 + 16 B), both streams once out (32 B) per observation, against
// ~10 passes over 12-byte key/value pairs for the two device-wide sorts; one launch instead of ~20.
constexpr int kWinIdxThreads = 512;

struct WinIdxSmem {   // layout of the dynamic shared memory (ints), for a window with <= max_np landmarks, <= 32 cameras
  int max_np, nchunk;
  __host__ __device__ size_t mask() const { return 0; }
  __host__ __device__ size_t sched() const { return mask() + max_np; }
  __host__ __device__ size_t pos_of() const { return sched() + max_np; }
  __host__ __device__ size_t ptr() const { return pos_of() + max_np; }                 // max_np + 1
  __host__ __device__ size_t bv() const { return ptr() + max_np + 1; }                 // 32 x nchunk
  __host__ __device__ size_t cp() const { return bv() + (size_t)32 * nchunk; }         // 32 x nchunk
  __host__ __device__ size_t wc() const { return cp() + (size_t)32 * nchunk; }         // (threads / 32) x 32
  __host__ __device__ size_t run() const { return wc() + (kWinIdxThreads / 32) * 32; } // 32 bucket cursors
  __host__ __device__ size_t cam() const { return run() + 32; }                        // 33 camera run starts
  __host__ __device__ size_t total() const { return cam() + 33 + 3; }
};

__global__ void __launch_bounds__(kWinIdxThreads) window_index_kernel(IndexIn in, IndexOut out, WinIdxSmem L) {
  extern __shared__ int sm[];
  unsigned* mask = reinterpret_cast<unsigned*>(sm + L.mask());
  int* sched = sm + L.sched();
  int* pos_of = sm + L.pos_of();
  int* ptr = sm + L.ptr();
  unsigned* bv = reinterpret_cast<unsigned*>(sm + L.bv());
  int* cp = sm + L.cp();
  int* wc = sm + L.wc();
  int* run = sm + L.run();
  int* cam = sm + L.cam();
  int* carry = cam + 33;
  constexpr int T = kWinIdxThreads, NW = T / 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int c0 = in.cam_off[b], nc = in.cam_off[b + 1] - c0;
  const int p0 = in.pt_off[b], np = in.pt_off[b + 1] - p0;
  const long long o0 = in.obs_off[b], o1 = in.obs_off[b + 1];
  const int nchunk = (np + 31) >> 5;

  for (int p = tid; p < np; p += T) mask[p] = 0u;
  if (tid < 32) run[tid] = 0;
  __syncthreads();
  // 1. camera mask per landmark; duplicate (camera, landmark) pairs and out-of-range indices are flagged
  for (long long i = o0 + tid; i < o1; i += T) {
    int c = in.obs_cam[i], p = in.obs_pt[i];
    if (c < 0 || c >= nc || p < 0 || p >= np) { out.flags[0] = 1; c = 0; p = 0; }
    if (np > 0) {
      const unsigned bit = 1u << c;
      if (atomicOr(&mask[p], bit) & bit) out.flags[1] = 1;
    }
  }
  __syncthreads();
  // 2. landmark schedule: stable counting sort by bucket 31 - min(observation count, 31) (landmark id order inside a
  //    bucket).  run[k]: landmarks of bucket k seen so far; first the bucket sizes, then their starts, then a pass in
  //    id order, T landmarks at a time: rank inside the warp by match_any, over the warps by a small table.
  for (int p = tid; p < np; p += T) atomicAdd(&run[31 - min(__popc(mask[p]), 31)], 1);
  __syncthreads();
  if (tid == 0) {
    int acc = 0;
    for (int k = 0; k < 32; ++k) { const int n = run[k]; run[k] = acc; acc += n; }
  }
  __syncthreads();
  for (int base = 0; base < np; base += T) {
    const int p = base + tid;
    const bool live = p < np;
    const int k = live ? 31 - min(__popc(mask[p]), 31) : 32 + lane;        // idle lanes match nobody
    const unsigned same = __match_any_sync(0xffffffffu, k);
    const int rank_w = __popc(same & ((1u << lane) - 1u));
    wc[warp * 32 + lane] = 0;
    __syncwarp();
    if (live && rank_w == 0) wc[warp * 32 + k] = __popc(same);            // one leader per key and warp
    __syncthreads();
    if (live) {
      int off = run[k] + rank_w;
      for (int w = 0; w < warp; ++w) off += wc[w * 32 + k];
      sched[off] = p;
      pos_of[p] = off;
    }
    __syncthreads();
    if (tid < 32) {
      int n = 0;
      for (int w = 0; w < NW; ++w) n += wc[w * 32 + tid];
      run[tid] += n;
    }
    __syncthreads();
  }
  // 3. pt_ptr: exclusive scan of the observation counts in schedule order
  if (tid == 0) carry[0] = 0;
  __syncthreads();
  for (int base = 0; base < np; base += T) {
    const int q = base + tid;
    const int v = q < np ? __popc(mask[sched[q]]) : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) wc[warp] = incl;
    __syncthreads();
    int before = carry[0];
    for (int w = 0; w < warp; ++w) before += wc[w];
    if (q < np) ptr[q] = before + incl - v;
    __syncthreads();
    if (tid == T - 1) carry[0] = before + incl;
    __syncthreads();
  }
  if (tid == 0) ptr[np] = carry[0];
  // 4. per camera: bit vector over the slots, observation count in front of every 32 slots, run starts
  for (int j = warp; j < nchunk; j += NW) {
    const int q = 32 * j + lane;
    const unsigned m = q < np ? mask[sched[q]] : 0u;
    for (int c = 0; c < nc; ++c) {
      const unsigned bits = __ballot_sync(0xffffffffu, (m >> c) & 1u);
      if (lane == c) bv[c * nchunk + j] = bits;
    }
  }
  __syncthreads();
  if (tid < nc) {
    int acc = 0;
    for (int j = 0; j < nchunk; ++j) { cp[tid * nchunk + j] = acc; acc += __popc(bv[tid * nchunk + j]); }
    cam[tid + 1] = acc;                    // observations of camera tid
  }
  __syncthreads();
  if (tid == 0) {
    cam[0] = 0;
    for (int c = 0; c < nc; ++c) cam[c + 1] += cam[c];
  }
  __syncthreads();
  // 5. offsets and the schedule out (absolute offsets; a window's last entry is its successor's first: same value)
  for (int q = tid; q <= np; q += T) out.pt_ptr[p0 + q] = (int)o0 + ptr[q];
  for (int c = tid; c <= nc; c += T) out.cam_ptr[c0 + c] = (int)o0 + cam[c];
  if (out.sched)
    for (int q = tid; q < np; q += T) out.sched[p0 + q] = sched[q];
  // 6. both streams
  if (np > 0) {
    for (long long i = o0 + tid; i < o1; i += T) {
      int c = in.obs_cam[i], p = in.obs_pt[i];
      if (c < 0 || c >= nc || p < 0 || p >= np) { c = 0; p = 0; }
      const float2 uv = in.obs_uv[i];
      const int s_ = pos_of[p];
      const unsigned m = mask[p];
      const long long a_pm = o0 + ptr[s_] + __popc(m & ((1u << c) - 1u));
      const long long a_cm = o0 + cam[c] + cp[c * nchunk + (s_ >> 5)] + __popc(bv[c * nchunk + (s_ >> 5)] & ((1u << (s_ & 31)) - 1u));
      ObsRec r; r.u = uv.x; r.v = uv.y; r.orig = (int)i;
      r.idx = c; out.pm[a_pm] = r;
      r.idx = s_; out.cm[a_cm] = r;
    }
  }
}

// the per-window builder applies when the streams are schedule-ordered (window path), every window has <= 32 cameras
// and its tables fit into shared memory
static bool window_index_fits(const IndexIn& in, const IndexOut& out, WinIdxSmem* L, size_t* bytes) {
  if (!out.sched || in.max_nc <= 0 || in.max_nc > 32 || in.max_np <= 0) return false;
  static const bool off = [] { const char* e = std::getenv("B2BA_INDEX_SORT"); return e && std::atoi(e) != 0; }();
  if (off) return false;                   // developer knob: force the global-sort builder (parity of the two builders)
  L->max_np = in.max_np; L->nchunk = (in.max_np + 31) / 32;
  *bytes = L->total() * sizeof(int);
  return *bytes <= 200 * 1024;
}

int build_index(const IndexIn& in, const IndexOut& out, cudaStream_t s) {
  cudaError_t e;
  e = cudaMemsetAsync(out.flags, 0, 2 * sizeof(int), s);
  if (e != cudaSuccess) return (int)e;
  const long long no = in.no;
  {
    WinIdxSmem L;
    size_t bytes = 0;
    if (window_index_fits(in, out, &L, &bytes)) {
      static std::mutex mtx;
      static bool configured[64] = {};
      int dev = 0;
      if ((e = cudaGetDevice(&dev)) != cudaSuccess) return (int)e;
      {
        std::lock_guard<std::mutex> lock(mtx);
        if (dev < 0 || dev >= 64 || !configured[dev]) {
          e = cudaFuncSetAttribute(window_index_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
          if (e != cudaSuccess) return (int)e;
          if (dev >= 0 && dev < 64) configured[dev] = true;
        }
      }
      window_index_kernel<<<in.nb, kWinIdxThreads, bytes, s>>>(in, out, L);
      return (int)cudaGetLastError();
    }
  }
  if (no == 0) {
    fill_ptr<<<(in.np_tot + 256) / 256, 256, 0, s>>>(out.pt_ptr, in.np_tot + 1, 0);
    fill_ptr<<<(in.nc_tot + 256) / 256, 256, 0, s>>>(out.cam_ptr, in.nc_tot + 1, 0);
    if (out.sched && in.np_tot > 0) return (int)sort_sched(in, out, nullptr, s);
    return (int)cudaGetLastError();
  }
  unsigned long long *k0 = nullptr, *k1 = nullptr;
  int *v0 = nullptr, *v1 = nullptr, *p1 = nullptr, *pos_of = nullptr;
  unsigned *ck0 = nullptr, *ck1 = nullptr;
  void* tmp = nullptr;
  size_t tmp_bytes = 0, t2 = 0;
  const int pbits = 32 + bits_for(in.np_tot), cbits = bits_for(in.nc_tot);
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, v1, no, 0, pbits, s);
  cub::DeviceRadixSort::SortPairs(nullptr, t2, ck0, ck1, v0, v1, no, 0, cbits, s);
  if (t2 > tmp_bytes) tmp_bytes = t2;
#define B2_TRY(x) do { e = (x); if (e != cudaSuccess) goto done; } while (0)
  B2_TRY(cudaMallocAsync(&k0, no * 8, s)); B2_TRY(cudaMallocAsync(&k1, no * 8, s));
  B2_TRY(cudaMallocAsync(&v0, no * 4, s)); B2_TRY(cudaMallocAsync(&v1, no * 4, s));
  B2_TRY(cudaMallocAsync(&p1, no * 4, s));
  B2_TRY(cudaMallocAsync(&ck0, no * 4, s)); B2_TRY(cudaMallocAsync(&ck1, no * 4, s));
  B2_TRY(cudaMallocAsync(&tmp, tmp_bytes, s));
  {
    const int nblk = (int)((no + 255) / 256);
    if (out.sched && in.np_tot > 0) {
      // window path: the landmarks of a window are walked by decreasing observation count; the point-major
      // stream (pm, pt_ptr) is laid out in that order so that a batch's records are one contiguous run
      B2_TRY(cudaMallocAsync(&pos_of, (size_t)in.np_tot * 4, s));
      B2_TRY(sort_sched(in, out, pos_of, s));
    }
    make_pm_keys<<<nblk, 256, 0, s>>>(in, pos_of, k0, v0, out.flags);
    B2_TRY(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, v1, no, 0, pbits, s));
    build_pm<<<nblk, 256, 0, s>>>(in, k1, v1, out.pm, out.pt_ptr, ck0, v0, out.flags);
    B2_TRY(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, ck0, ck1, v0, p1, no, 0, cbits, s));
    build_cm<<<nblk, 256, 0, s>>>(in, ck1, p1, out.pm, k1, pos_of ? out.sched : nullptr, out.cm, out.cam_ptr);
    B2_TRY(cudaGetLastError());
  }
done:
#undef B2_TRY
  cudaFreeAsync(k0, s); cudaFreeAsync(k1, s); cudaFreeAsync(v0, s); cudaFreeAsync(v1, s);
  cudaFreeAsync(p1, s); cudaFreeAsync(ck0, s); cudaFreeAsync(ck1, s); cudaFreeAsync(tmp, s); cudaFreeAsync(pos_of, s);
  return (int)e;
}

namespace {
__global__ void obs_terms_kernel(IndexIn in, Params P, const double* __restrict__ poses,
                                 const double* __restrict__ pts, double* r, double* w, double* jc,
                                 double* jp) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= in.no) return;
  const int b = find_window(in.obs_off, in.nb, i);
  const double* pose = poses + 12 * (size_t)(in.cam_off[b] + in.obs_cam[i]);
  const double* X = pts + 3 * (size_t)(in.pt_off[b] + in.obs_pt[i]);
  const float2 uv = in.obs_uv[i];
  Proj o;
  project(P, pose, X[0], X[1], X[2], uv.x, uv.y, o);
  double a[12], c[6];
  jac_cam(P, o, a);
  jac_point(P, pose, o, c);
  if (r) { r[2 * i] = o.r0; r[2 * i + 1] = o.r1; }
  if (w) w[i] = o.w;
  if (jc) for (int k = 0; k < 12; ++k) jc[12 * i + k] = a[k];
  if (jp) for (int k = 0; k < 6; ++k) jp[6 * i + k] = c[k];
}
}  // namespace

namespace {
// reprojection error in pixels per observation, ORIGINAL observation order; -1 behind the cheirality gate
__global__ void obs_error_kernel(IndexIn in, Params P, const double* __restrict__ poses,
                                 const double* __restrict__ pts, double* err) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= in.no) return;
  const int b = find_window(in.obs_off, in.nb, i);
  const double* pose = poses + 12 * (size_t)(in.cam_off[b] + in.obs_cam[i]);
  const double* X = pts + 3 * (size_t)(in.pt_off[b] + in.obs_pt[i]);
  const float2 uv = in.obs_uv[i];
  Params Q = P;
  Q.loss = 0;                        // the plain pixel residual, whatever the solver's loss
  Proj o;
  project(Q, pose, X[0], X[1], X[2], uv.x, uv.y, o);
  err[i] = o.valid ? sqrt(o.r0 * o.r0 + o.r1 * o.r1) : -1.0;
}
}  // namespace

int obs_errors(const IndexIn& in, const Params& P, const double* poses, const double* pts, double* err,
               cudaStream_t s) {
  if (in.no == 0) return 0;
  obs_error_kernel<<<(int)((in.no + 255) / 256), 256, 0, s>>>(in, P, poses, pts, err);
  return (int)cudaGetLastError();
}

int dump_obs_terms(const IndexIn& in, const Params& P, const double* poses, const double* pts,
                   double* r, double* w, double* jc, double* jp, cudaStream_t s) {
  if (in.no == 0) return 0;
  obs_terms_kernel<<<(int)((in.no + 127) / 128), 128, 0, s>>>(in, P, poses, pts, r, w, jc, jp);
  return (int)cudaGetLastError();
}

}  // namespace b2
