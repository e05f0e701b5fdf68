// Probe for the C5 multi-GPU exchange design: one process per GPU (fork), peer buffers mapped with CUDA IPC,
// exchanges done INSIDE a persistent kernel by stores into peer memory + epoch flags (no NCCL, no host).
//   test A: one-shot all-gather of a slice (each rank stores its slice into every peer's buffer, then a flag),
//   test B: one-shot all-reduce (each rank stores its whole partial vector into slot[rank] of every peer; every
//           rank sums the N slots in rank order -> deterministic, identical on all ranks),
// both repeated `iters` times back to back in ONE launch; reports us per exchange and checks the data.
// usage: ipc_probe <nranks> [slice_doubles=768] [vec_doubles=6000] [iters=2000]
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ipc_probe tools/ipc_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "rank %d: %s -> %s\n", rank, #x, cudaGetErrorString(e_)); _exit(2); } } while (0)

constexpr int kMaxRanks = 8;
struct Shared {
  cudaIpcMemHandle_t h[kMaxRanks];
  volatile int arrive[256];
  double us[kMaxRanks][4], us2[kMaxRanks][4];
  int ok[kMaxRanks];
};

static void host_barrier(Shared* sh, int nranks, int& epoch) {
  const int slot = epoch % 256;
  __sync_fetch_and_add(&sh->arrive[slot], 1);
  while (sh->arrive[slot] < nranks) usleep(50);
  ++epoch;
}

struct Peers { double* buf[kMaxRanks]; };

// layout of every rank's buffer (doubles): [2 parities][nranks slots][vec] data, then flags (as long long) [2][nranks]
__device__ __forceinline__ void st_flag(long long* p, long long v) {
  asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ long long ld_flag(const long long* p) {
  long long v;
  asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// MODE 0: all-gather of `cnt` doubles per rank; MODE 1: all-reduce of `cnt` doubles.  One CTA.
template <int MODE>
__global__ void __launch_bounds__(1024) exch_kernel(Peers P, int rank, int nranks, int cnt, int vec, int iters, double* result, int fence_all) {
  const int tid = threadIdx.x;
  const size_t slot_doubles = (size_t)vec;
  const size_t data_doubles = 2 * (size_t)nranks * slot_doubles;
  double* mine = P.buf[rank];
  long long* myflags = reinterpret_cast<long long*>(mine + data_doubles);
  double acc = 0.0;
  int bad = 0;
  for (int it = 1; it <= iters; ++it) {
    const int par = it & 1;
    // produce: value depends on (rank, it, i) so stale data is detected
    for (int r = 0; r < nranks; ++r) {
      double* dst = P.buf[r] + ((size_t)par * nranks + rank) * slot_doubles;
      for (int i = tid; i < cnt; i += blockDim.x) dst[i] = (double)(rank + 1) * 1e-3 * it + i;
    }
    if (fence_all) __threadfence_system();     // every writer fences its own stores
    __syncthreads();
    if (tid < nranks) {
      if (!fence_all) __threadfence_system();  // one fence after the barrier (cumulativity)
      st_flag(reinterpret_cast<long long*>(P.buf[tid] + data_doubles) + par * nranks + rank, it);
    }
    if (tid < nranks) { while (ld_flag(myflags + par * nranks + tid) < it) {} }
    __syncthreads();
    // consume (L2 loads: peers wrote the lines, L1 may hold the copy of two iterations ago) and verify exactly
    const double* base = mine + (size_t)par * nranks * slot_doubles;
    for (int i = tid; i < cnt; i += blockDim.x) {
      double v = 0.0;
      for (int r = 0; r < nranks; ++r) {
        const double x = __ldcg(base + (size_t)r * slot_doubles + i);
        if (fabs(x - ((double)(r + 1) * 1e-3 * it + i)) > 1e-7) ++bad;   // a stale value (iteration it - 2) is off by >= 2e-3
        v += x;
      }
      acc += v;
    }
    __syncthreads();
  }
  __shared__ int sbad;
  if (tid == 0) sbad = 0;
  __syncthreads();
  if (bad) atomicAdd(&sbad, bad);
  __syncthreads();
  if (tid == 0) { result[0] = acc; result[1] = (double)sbad; }
}

int main(int argc, char** argv) {
  const int nranks = argc > 1 ? atoi(argv[1]) : 2;
  const int slice = argc > 2 ? atoi(argv[2]) : 768, vec = argc > 3 ? atoi(argv[3]) : 6000, iters = argc > 4 ? atoi(argv[4]) : 2000;
  Shared* sh = (Shared*)mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
  memset(sh, 0, sizeof(Shared));
  int rank = 0;
  pid_t kids[kMaxRanks];
  for (int r = 1; r < nranks; ++r) {
    pid_t p = fork();
    if (p == 0) { rank = r; break; }
    kids[r] = p;
  }
  int epoch = 0;
  CK(cudaSetDevice(rank));
  const size_t doubles = 2 * (size_t)nranks * vec + 2 * nranks + 16;
  double* buf = nullptr;
  CK(cudaMalloc(&buf, doubles * sizeof(double)));
  CK(cudaMemset(buf, 0, doubles * sizeof(double)));
  CK(cudaDeviceSynchronize());
  CK(cudaIpcGetMemHandle(&sh->h[rank], buf));
  host_barrier(sh, nranks, epoch);
  Peers P;
  for (int r = 0; r < nranks; ++r) {
    if (r == rank) { P.buf[r] = buf; continue; }
    void* p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, sh->h[r], cudaIpcMemLazyEnablePeerAccess));
    P.buf[r] = (double*)p;
  }
  double* d_res; CK(cudaMalloc(&d_res, 16));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  int ok = 1;
  for (int mode = 0; mode < 2; ++mode) {
    const int cnt = mode == 0 ? slice : vec;
    for (int cfg = 0; cfg < 2; ++cfg) {
      const int threads = cfg == 0 ? 256 : 1024;
      for (int fence_all = 0; fence_all < 2; ++fence_all) {
        CK(cudaMemset(buf, 0, doubles * sizeof(double)));
        CK(cudaDeviceSynchronize());
        host_barrier(sh, nranks, epoch);
        CK(cudaEventRecord(e0));
        if (mode == 0) exch_kernel<0><<<1, threads>>>(P, rank, nranks, cnt, vec, iters, d_res, fence_all);
        else exch_kernel<1><<<1, threads>>>(P, rank, nranks, cnt, vec, iters, d_res, fence_all);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        double res[2]; CK(cudaMemcpy(res, d_res, 16, cudaMemcpyDeviceToHost));
        if (res[1] != 0.0) { ok = 0; fprintf(stderr, "rank %d mode %d threads %d fence_all %d: %.0f stale values\n", rank, mode, threads, fence_all, res[1]); }
        if (fence_all == 0) sh->us[rank][mode * 2 + cfg] = ms * 1e3 / iters;
        else sh->us2[rank][mode * 2 + cfg] = ms * 1e3 / iters;
        host_barrier(sh, nranks, epoch);
      }
    }
  }
  sh->ok[rank] = ok;
  host_barrier(sh, nranks, epoch);
  if (rank == 0) {
    double mx[4] = {0, 0, 0, 0}, mx2[4] = {0, 0, 0, 0};
    int all_ok = 1;
    for (int r = 0; r < nranks; ++r) {
      all_ok &= sh->ok[r];
      for (int k = 0; k < 4; ++k) { if (sh->us[r][k] > mx[k]) mx[k] = sh->us[r][k]; if (sh->us2[r][k] > mx2[k]) mx2[k] = sh->us2[r][k]; }
    }
    printf("{\"nranks\": %d, \"allgather_doubles_per_rank\": %d, \"allgather_us_256thr\": %.2f, \"allgather_us_1024thr\": %.2f, "
           "\"allreduce_doubles\": %d, \"allreduce_us_256thr\": %.2f, \"allreduce_us_1024thr\": %.2f, "
           "\"fence_by_every_writer\": {\"allgather_us_256thr\": %.2f, \"allgather_us_1024thr\": %.2f, \"allreduce_us_256thr\": %.2f, \"allreduce_us_1024thr\": %.2f}, "
           "\"data_ok\": %s}\n",
           nranks, slice, mx[0], mx[1], vec, mx[2], mx[3], mx2[0], mx2[1], mx2[2], mx2[3], all_ok ? "true" : "false");
    for (int r = 1; r < nranks; ++r) { int st; waitpid(kids[r], &st, 0); }
  }
  for (int r = 0; r < nranks; ++r) if (r != rank) cudaIpcCloseMemHandle(P.buf[r]);
  return 0;
}
