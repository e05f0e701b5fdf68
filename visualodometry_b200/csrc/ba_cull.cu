// outlier culling after bundle adjustment, on the device, in one kernel.
//
// The reference gates a landmark on its reprojection error (`frontend.py:139-142`: |projectPoints(X) - uv| <
// max_reproj_err, `config.py:32`; depth gate `config.py:31`) and marks an untracked keypoint with id -1 (`vo.py:53`);
// its map pruning (`vo.py:35-47`) drops landmarks by age.  After BA the same gate is applied to every observation of
// the window at the refined state: cull_kernel evaluates the error, decides keep, compacts the observation arrays in
// the caller's order (stable), and counts what survives per landmark and per window.
//
// Single pass over the observations: a CTA takes a tile of 2048 consecutive observations (8 per thread, blocked, so the
// order inside a tile is the original one), scans its keep flags, and obtains the number of kept observations in front
// of its tile by decoupled look-back over the tile descriptors (status | count packed in one 64-bit word, tiles handed
// out by an atomic ticket so that every predecessor of a running tile is running or done).  HBM traffic = the
// observation arrays once in, the kept part once out: 16 B + 16 B x kept fraction per observation.
#include "ba_index.cuh"

namespace b2 {

namespace {

constexpr int kCullThreads = 256, kCullIpt = 8, kCullTile = kCullThreads * kCullIpt;
constexpr unsigned long long kAggregate = 1ull << 62, kPrefix = 2ull << 62, kValueMask = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_acquire(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// work: [0] ticket counter, [1 ...] one descriptor per tile (zeroed before the launch)
__global__ void __launch_bounds__(kCullThreads) cull_kernel(IndexIn in, Params P, const double* __restrict__ poses,
                                                            const double* __restrict__ pts, double max_err,
                                                            unsigned char* keep, int* live, long long* win_kept,
                                                            int* out_cam, int* out_pt, float2* out_uv, int* kept_index,
                                                            long long* n_kept, unsigned long long* work) {
  __shared__ long long s_tile;
  __shared__ long long s_prefix;
  __shared__ int s_warp[kCullThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = (long long)atomicAdd(work, 1ull);
  __syncthreads();
  const long long tile = s_tile;
  const long long first = tile * kCullTile + (long long)tid * kCullIpt;
  P.loss = 0;                          // the plain pixel residual, whatever the solver's loss
  unsigned flags = 0;
  int cam[kCullIpt], pt[kCullIpt];
  float2 uv[kCullIpt];
  int wb = -1;                         // window of the previous item of this thread
  long long wcount = 0;
#pragma unroll
  for (int k = 0; k < kCullIpt; ++k) {
    const long long i = first + k;
    if (i >= in.no) break;
    const int b = find_window(in.obs_off, in.nb, i);
    cam[k] = in.obs_cam[i]; pt[k] = in.obs_pt[i]; uv[k] = in.obs_uv[i];
    const int g = in.pt_off[b] + pt[k];
    const double* pose = poses + 12 * (size_t)(in.cam_off[b] + cam[k]);
    const double* X = pts + 3 * (size_t)g;
    Proj o;
    project(P, pose, X[0], X[1], X[2], uv[k].x, uv[k].y, o);
    const double err = o.valid ? sqrt(o.r0 * o.r0 + o.r1 * o.r1) : -1.0;
    const bool kp = !(err < 0.0) && !(err > max_err);     // the complement of the host rule (err < 0) | (err > max_err)
    if (keep) keep[i] = kp ? 1 : 0;
    if (kp) {
      flags |= 1u << k;
      if (live) atomicAdd(live + g, 1);
    }
    if (win_kept) {
      if (b != wb) {
        if (wcount) atomicAdd(reinterpret_cast<unsigned long long*>(win_kept + wb), (unsigned long long)wcount);
        wb = b; wcount = 0;
      }
      wcount += kp ? 1 : 0;
    }
  }
  if (win_kept && wcount) atomicAdd(reinterpret_cast<unsigned long long*>(win_kept + wb), (unsigned long long)wcount);
  // exclusive scan of the per-thread counts over the CTA
  const int cnt = __popc(flags);
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kCullThreads / 32; ++w) {
    if (w < warp) before += s_warp[w];
    total += s_warp[w];
  }
  const int excl = before + incl - cnt;
  // decoupled look-back: publish this tile's count, add up the predecessors' until one carries an inclusive prefix
  if (tid == 0) {
    unsigned long long* desc = work + 1;
    long long prefix = 0;
    if (tile == 0) {
      st_release(desc, kPrefix | (unsigned long long)total);
    } else {
      st_release(desc + tile, kAggregate | (unsigned long long)total);
      for (long long j = tile - 1; j >= 0; --j) {
        unsigned long long v;
        while (((v = ld_acquire(desc + j)) >> 62) == 0) __nanosleep(20);
        prefix += (long long)(v & kValueMask);
        if ((v >> 62) == 2) break;
      }
      st_release(desc + tile, kPrefix | (unsigned long long)(prefix + total));
    }
    s_prefix = prefix;
    if (n_kept && (tile + 1) * (long long)kCullTile >= in.no) *n_kept = prefix + total;   // the last tile
  }
  __syncthreads();
  long long pos = s_prefix + excl;
#pragma unroll
  for (int k = 0; k < kCullIpt; ++k) {
    if (flags & (1u << k)) {
      if (out_cam) out_cam[pos] = cam[k];
      if (out_pt) out_pt[pos] = pt[k];
      if (out_uv) out_uv[pos] = uv[k];
      if (kept_index) kept_index[pos] = (int)(first + k);
      ++pos;
    }
  }
}

}  // namespace

size_t cull_work_bytes(long long no) { return (size_t)(1 + (no + kCullTile - 1) / kCullTile) * sizeof(unsigned long long); }

int cull_observations(const IndexIn& in, const Params& P, const double* poses, const double* pts, double max_err,
                      unsigned char* keep, int* live, long long* win_kept, int* out_cam, int* out_pt, float2* out_uv,
                      int* kept_index, long long* n_kept, void* work, cudaStream_t s) {
  cudaError_t e;
  if (live && (e = cudaMemsetAsync(live, 0, (size_t)in.np_tot * sizeof(int), s)) != cudaSuccess) return (int)e;
  if (win_kept && (e = cudaMemsetAsync(win_kept, 0, (size_t)in.nb * sizeof(long long), s)) != cudaSuccess) return (int)e;
  if (n_kept && (e = cudaMemsetAsync(n_kept, 0, sizeof(long long), s)) != cudaSuccess) return (int)e;
  if (in.no == 0) return 0;
  if ((e = cudaMemsetAsync(work, 0, cull_work_bytes(in.no), s)) != cudaSuccess) return (int)e;
  const long long tiles = (in.no + kCullTile - 1) / kCullTile;
  cull_kernel<<<(unsigned)tiles, kCullThreads, 0, s>>>(in, P, poses, pts, max_err, keep, live, win_kept, out_cam, out_pt,
                                                       out_uv, kept_index, n_kept, static_cast<unsigned long long*>(work));
  return (int)cudaGetLastError();
}

}  // namespace b2
