// C ABI of the b2ba engine (include/b2ba.h): handle, problem upload, solve, read-back, dumps.
// All compute is CUDA (sm_100a); there is NO CPU fallback -- every entry point that needs the
// device fails with B2BA_E_CUDA when it is missing.
#include "../../include/b2ba.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <dlfcn.h>
#include <mutex>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "ba_index.cuh"
#include "devbuf.cuh"
#include "ba_large.cuh"
#include "ba_window.cuh"
#include "ba_pose.cuh"

using namespace b2;

namespace {

enum Path { PATH_NONE = 0, PATH_WIN_SMALL, PATH_WIN_BIG, PATH_LARGE };

}  // namespace

struct b2ba {
  int device = 0, sms = 148;
  b2ba_options opt;
  std::string err;
  int B = 0, nc_tot = 0, np_tot = 0;
  long long no = 0;
  int max_nfree = 0, max_nc = 0;
  Path path = PATH_NONE;
  double intr[4] = {0, 0, 0, 0};
  std::vector<int> h_cam_off, h_pt_off;
  std::vector<long long> h_obs_off;
  std::vector<unsigned char> h_fixed;
  DevBuf<int> d_cam_off, d_pt_off, d_obs_cam, d_obs_pt, d_pt_ptr, d_cam_ptr, d_flags, d_hist_accept, d_sched;
  DevBuf<long long> d_obs_off;
  DevBuf<float2> d_uv;
  DevBuf<unsigned char> d_fixed;
  DevBuf<double> d_poses, d_poses0, d_pts0, d_pts1, d_ptsinit, d_scratch, d_hist_cost, d_hist_mu, d_dbg;
  DevBuf<ObsRec> d_pm, d_cm;
  DevBuf<b2ba_summary> d_summ;
  // large path
  LargeIndex lix;
  DevBuf<ObsRec> l_cm;
  DevBuf<int> l_cam_ptr;
  DevBuf<double> l_poses_tri, l_lin_part, l_Dc, l_Sg, l_dc, l_r, l_z, l_q, l_Minv, l_part, l_sums, l_band, l_binv, l_band_ws, l_zbuf, l_dense, l_dense_winv, l_dense_x, l_tbuf;
  DevBuf<int> l_dense_flags;
  int l_dense_ld = 0;
  // motion-only BA (b2ba_refine_poses): own buffers, independent of the problem set with b2ba_set_problem
  DevBuf<long long> r_off;
  DevBuf<double> r_poses, r_pts, r_hist_cost, r_hist_mu;
  DevBuf<float2> r_uv;
  DevBuf<b2ba_summary> r_summ;
  DevBuf<int> r_hist_accept, r_cnt, q_idx, q_score, q_best;
  DevBuf<double> q_pts, q_poses;
  DevBuf<float2> q_uv;
  DevBuf<long long> q_off;
  DevBuf<b2ba_summary> q_summ;
  DevBuf<unsigned char> r_mask;
  int r_nb = 0;
  bool sched_order = false;   // pm / pt_ptr laid out in landmark-schedule order (window path), else canonical
  int l_band_P = 1;  // interiors of the substructured banded solve (1 = sequential)
  int l_bw = -1;     // block half-bandwidth of S when the banded Cholesky is selected, else -1
  DevBuf<LargeCtrl> l_ctrl;
  int p_lo = 0, p_hi = 0;
  bool range_set = false;
  cudaStream_t last_stream = nullptr;
  bool solved = false;
  // timings (options.flags bit 4)
  std::vector<cudaEvent_t> t_ev;       // 7 per LM iteration (large path) or 2 (window path)
  cudaEvent_t t_setup[2] = {nullptr, nullptr};
  int t_iters = 0;
  bool t_valid = false, t_setup_valid = false;
  // multi-GPU
  int rank = 0, nranks = 1;
  void* nccl_comm = nullptr;
  DevBuf<double> l_arena, l_Sl;        // peer-exchange arena (mapped into every rank), owned rows of S
  void* peer_arena[kMaxRanks] = {};    // every rank's arena as mapped into this process
  Peer peer = {};
  // pipelined batch call (b2ba_optimize_batch): child engines (one per chunk) and their streams / events
  std::vector<b2ba*> kids;
  cudaStream_t pipe_copy = nullptr, pipe_solve = nullptr, pipe_solve2 = nullptr, pipe_back = nullptr;
  std::vector<cudaEvent_t> ev_in, ev_done;
  int* pinned_flags = nullptr;     // 2 ints per chunk (index-builder range / duplicate flags)
  char* pinned_stage = nullptr;    // summaries + rebased offsets (pinned: async copies never block the enqueue loop)
  size_t pinned_stage_bytes = 0;
  int wave_small = 0, wave_big = 0;   // windows one launch wave holds (CTAs per SM x SMs)
};

namespace {

int fail(b2ba* h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}
int cuda_fail(b2ba* h, cudaError_t e, const char* where) {
  return fail(h, B2BA_E_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
#define CU(x, where) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return cuda_fail(h, e_, where); } while (0)

// Every entry point runs on the handle's device and gives the caller's current device back on return (a torch
// process keeps its own notion of the current device; the engine must not change it behind its back).
struct DeviceGuard {
  int prev = -1, dev;
  cudaError_t err;
  explicit DeviceGuard(int d) : dev(d) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
  }
  ~DeviceGuard() { if (prev >= 0 && prev != dev) cudaSetDevice(prev); }
};

Params make_params(const b2ba* h, int max_iters) {
  Params P;
  P.fx = h->intr[0]; P.fy = h->intr[1]; P.cx = h->intr[2]; P.cy = h->intr[3];
  P.z_min = h->opt.z_min; P.huber_a = h->opt.loss_scale; P.init_mu = h->opt.init_mu;
  P.ftol = h->opt.ftol; P.gtol = h->opt.gtol; P.ptol = h->opt.ptol;
  P.loss = h->opt.loss; P.max_iters = max_iters > 0 ? max_iters : h->opt.max_iters;
  return P;
}

IndexIn make_index_in(const b2ba* h) {
  IndexIn in;
  in.nb = h->B; in.nc_tot = h->nc_tot; in.np_tot = h->np_tot; in.no = h->no;
  in.cam_off = h->d_cam_off.p; in.pt_off = h->d_pt_off.p; in.obs_off = h->d_obs_off.p;
  in.obs_cam = h->d_obs_cam.p; in.obs_pt = h->d_obs_pt.p; in.obs_uv = h->d_uv.p;
  if ((int)h->h_pt_off.size() == h->B + 1 && (int)h->h_cam_off.size() == h->B + 1) {
    for (int b = 0; b < h->B; ++b) {
      in.max_np = std::max(in.max_np, h->h_pt_off[b + 1] - h->h_pt_off[b]);
      in.max_nc = std::max(in.max_nc, h->h_cam_off[b + 1] - h->h_cam_off[b]);
    }
  }
  return in;
}

// window path: landmark rows between the caller's order and the engine's schedule order (src != dst)
int permute_pts(b2ba* h, const double* src, double* dst, bool to_sched, cudaStream_t s) {
  return permute_points(h->B, h->np_tot, h->d_pt_off.p, h->d_sched.p, src, dst, to_sched, s);
}

WinData make_win(b2ba* h, bool hist, double* dbg) {
  WinData d;
  d.nb = h->B; d.cam_off = h->d_cam_off.p; d.pt_off = h->d_pt_off.p;
  d.pm = h->d_pm.p; d.pt_ptr = h->d_pt_ptr.p; d.cm = h->d_cm.p; d.cam_ptr = h->d_cam_ptr.p; d.sched = h->d_sched.p;
  d.fixed = h->d_fixed.p; d.poses = h->d_poses.p; d.pts[0] = h->d_pts0.p; d.pts[1] = h->d_pts1.p;
  d.pt_scratch = h->d_scratch.p; d.summaries = h->d_summ.p;
  d.hist_cost = hist ? h->d_hist_cost.p : nullptr;
  d.hist_mu = hist ? h->d_hist_mu.p : nullptr;
  d.hist_accept = hist ? h->d_hist_accept.p : nullptr;
  d.dbg = dbg;
  return d;
}

// few windows: the latency path, one 8/16-CTA cluster per window (B2BA_CLUSTER=0 or options.flags bit 1 force
// one CTA per window)
int launch_window_path(const b2ba* h, const WinData& wd, const Params& P, cudaStream_t s) {
  static const bool cluster_on = [] { const char* e = std::getenv("B2BA_CLUSTER"); return !e || std::atoi(e) != 0; }();
  const bool clus = cluster_on && !(h->opt.flags & 2) && h->B <= 18;
  const int variant = ((h->opt.flags & 4) ? kWinSmall256 : 0) | ((h->opt.flags & 8) ? kWinCluster8 : 0);
  if (h->path == PATH_WIN_SMALL) return clus ? launch_window_small_cluster(wd, P, s, variant) : launch_window_small(wd, P, s, variant);
  return clus ? launch_window_big_cluster(wd, P, s, variant) : launch_window_big(wd, P, s);
}

cudaMemcpyKind in_kind(int mem) { return mem == B2BA_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice; }
cudaMemcpyKind out_kind(int mem) { return mem == B2BA_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost; }


// ---------------------------------------------------------------------------------------------
// NCCL through dlopen: no link-time dependency, single-GPU use never touches it.  In a torch
// process the already-loaded libnccl.so.2 (torch's bundled 2.28.x) is the one that resolves.
struct NcclApi {
  typedef struct { char internal[128]; } UniqueId;
  int (*GetUniqueId)(UniqueId*) = nullptr;
  int (*CommInitRank)(void**, int, UniqueId, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};
NcclApi& nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (lib) {
      api.GetUniqueId = (int (*)(NcclApi::UniqueId*))dlsym(lib, "ncclGetUniqueId");
      api.CommInitRank = (int (*)(void**, int, NcclApi::UniqueId, int))dlsym(lib, "ncclCommInitRank");
      api.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(lib, "ncclAllReduce");
      api.CommDestroy = (int (*)(void*))dlsym(lib, "ncclCommDestroy");
      api.GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
      api.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(lib, "ncclAllGather");
      api.ok = api.GetUniqueId && api.CommInitRank && api.AllReduce && api.AllGather && api.CommDestroy;
    }
  }
  return api;
}
// Bootstrap collectives (set-up only; the LM loop exchanges through peer memory, see ba_large.cuh):
// all-gather of `bytes` per rank (device buffers) and a barrier.
int nccl_allgather_bytes(b2ba* h, const void* send, void* recv, size_t bytes) {
  int rc = nccl_api().AllGather(send, recv, bytes, /*ncclChar*/ 0, h->nccl_comm, nullptr);
  if (rc == 0 && cudaStreamSynchronize(nullptr) != cudaSuccess) rc = -1;
  if (rc != 0) {
    h->err = std::string("ncclAllGather: ") + (rc > 0 && nccl_api().GetErrorString ? nccl_api().GetErrorString(rc) : "error");
    return B2BA_E_NCCL;
  }
  return 0;
}
int nccl_barrier(b2ba* h) {
  DevBuf<double> one;
  if (one.alloc(1) != cudaSuccess) return fail(h, B2BA_E_CUDA, "alloc");
  cudaMemset(one.p, 0, sizeof(double));
  int rc = nccl_api().AllReduce(one.p, one.p, 1, /*ncclFloat64*/ 8, /*ncclSum*/ 0, h->nccl_comm, nullptr);
  if (rc == 0 && cudaStreamSynchronize(nullptr) != cudaSuccess) rc = -1;
  if (rc != 0) return fail(h, B2BA_E_NCCL, "ncclAllReduce (barrier) failed");
  return 0;
}

// unmap the peers' arenas (own entry is not a mapping)
void close_peers(b2ba* h) {
  for (int r = 0; r < kMaxRanks; ++r) {
    if (h->peer_arena[r] && r != h->rank) cudaIpcCloseMemHandle(h->peer_arena[r]);
    h->peer_arena[r] = nullptr;
  }
}

// nranks > 1: one exchange arena per rank, same layout everywhere, mapped into every rank with CUDA IPC
int setup_arena(b2ba* h) {
  const size_t n = 6 * (size_t)h->lix.nfree, nblk = h->lix.nblk, np = h->np_tot;
  auto up = [](size_t v) { return (v + 15) / 16 * 16; };   // 128-byte alignment of every region
  Peer& x = h->peer;
  size_t off = 0;
  x.off_flags = off; off += up(kMaxRanks);
  x.off_sums = off; off += up(kMaxRanks * 8);
  x.off_qx = off; off += up(2 * n);
  x.off_minv = off; off += up((size_t)h->lix.nfree * 37);
  x.off_r = off; off += up(n);
  x.off_sg = off; off += up(nblk * 36 + n);
  x.off_pts_tri = off; off += up(np * 3);
  x.rank = h->rank; x.nranks = h->nranks;
  if (off > h->l_arena.n || !h->l_arena.p) {
    // (re)allocation: every rank takes this branch together (the layout depends on the problem only)
    if (h->l_arena.p) {
      for (cudaEvent_t e : h->t_ev) cudaEventDestroy(e);
  if (h->t_setup[0]) { cudaEventDestroy(h->t_setup[0]); cudaEventDestroy(h->t_setup[1]); }
  close_peers(h);
      int rc = nccl_barrier(h);                    // nobody still maps the old arena
      if (rc) return rc;
    }
    CU(h->l_arena.alloc(off), "alloc arena");
    CU(cudaMemset(h->l_arena.p, 0, off * sizeof(double)), "memset arena");
    CU(cudaDeviceSynchronize(), "arena");
    cudaIpcMemHandle_t mine;
    CU(cudaIpcGetMemHandle(&mine, h->l_arena.p), "cudaIpcGetMemHandle");
    DevBuf<cudaIpcMemHandle_t> snd, rcv;
    CU(snd.alloc(1), "alloc"); CU(rcv.alloc(h->nranks), "alloc");
    CU(cudaMemcpy(snd.p, &mine, sizeof(mine), cudaMemcpyHostToDevice), "copy handle");
    int rc = nccl_allgather_bytes(h, snd.p, rcv.p, sizeof(mine));
    if (rc) return rc;
    std::vector<cudaIpcMemHandle_t> all(h->nranks);
    CU(cudaMemcpy(all.data(), rcv.p, sizeof(mine) * h->nranks, cudaMemcpyDeviceToHost), "read handles");
    for (int r = 0; r < h->nranks; ++r) {
      if (r == h->rank) { h->peer_arena[r] = h->l_arena.p; continue; }
      void* ptr = nullptr;
      const cudaError_t e = cudaIpcOpenMemHandle(&ptr, all[r], cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) return cuda_fail(h, e, "cudaIpcOpenMemHandle (peer arenas need peer access between the GPUs of one box)");
      h->peer_arena[r] = ptr;
    }
    rc = nccl_barrier(h);                          // every arena is zeroed and mapped before anyone stores into it
    if (rc) return rc;
  }
  for (int r = 0; r < kMaxRanks; ++r) x.arena[r] = (double*)h->peer_arena[r];
  return B2BA_OK;
}

LargeData make_large(b2ba* h, bool hist) {
  LargeData d;
  d.nc = h->nc_tot; d.np = h->np_tot; d.nfree = h->lix.nfree; d.n = 6 * h->lix.nfree;
  d.p_lo = h->p_lo; d.p_hi = h->p_hi; d.rank = h->rank; d.nranks = h->nranks;
  d.sms = h->sms;
  d.peer = h->peer;
  d.f_lo = h->lix.f_lo; d.f_hi = h->lix.f_hi; d.nlocal = h->lix.nlocal;
  d.lrow_ptr = h->lix.lrow_ptr.p; d.lcol = h->lix.lcol.p; d.lsrc = h->lix.lsrc.p; d.ldiag = h->lix.ldiag.p;
  d.Sl = h->l_Sl.p;
  d.pm = h->d_pm.p; d.pt_ptr = h->d_pt_ptr.p; d.cm = h->l_cm.p; d.cam_ptr = h->l_cam_ptr.p;
  d.fidx = h->lix.fidx.p; d.blk_ptr = h->lix.blk_ptr.p; d.blk_col = h->lix.blk_col.p; d.nblk = h->lix.nblk;
  d.frow_ptr = h->lix.frow_ptr.p; d.fcol = h->lix.fcol.p; d.fsrc = h->lix.fsrc.p;
  d.poses = h->d_poses.p; d.poses_tri = h->l_poses_tri.p; d.pts = h->d_pts0.p; d.pts_tri = h->d_pts1.p;
  d.lin_part = h->l_lin_part.p; d.lin = h->l_lin_part.p;     // the camera pass is replicated on nranks > 1
  d.Dc = h->l_Dc.p; d.pt_scratch = h->d_scratch.p; d.Sg = h->l_Sg.p; d.dc = h->l_dc.p;
  d.r = h->l_r.p; d.z = h->l_z.p; d.q = h->l_q.p; d.Minv = h->l_Minv.p; d.tbuf = h->l_tbuf.p; d.part = h->l_part.p;
  d.pair_ptr = h->lix.pair_ptr.p; d.pair_ab = h->lix.pair_ab.p; d.blk_row = h->lix.blk_row.p; d.zbuf = h->l_zbuf.p;
  d.pm_pt = h->lix.pm_pt.p; d.a_lo = h->lix.a_lo; d.a_hi = h->lix.a_hi;
  {
    // average pair-list length of the off-diagonal blocks (a diagonal list = the camera's observations)
    const long long noff = h->lix.nblk - h->lix.nfree;
    const long long per = noff > 0 ? (h->lix.npairs - (h->lix.a_hi - h->lix.a_lo)) / noff : 1000;
    d.pair_lanes = per < 10 ? 1 : per < 48 ? 8 : 32;
  }
  d.bw = h->l_bw; d.band = h->l_band.p; d.binv = h->l_binv.p; d.band_P = h->l_band_P; d.band_ws = h->l_band_ws.p;
  d.dense = h->l_dense_ld ? h->l_dense.p : nullptr; d.dense_ld = h->l_dense_ld; d.dense_winv = h->l_dense_winv.p;
  d.dense_x = h->l_dense_x.p; d.dense_flags = h->l_dense_flags.p;
  d.sums_part = h->l_sums.p; d.ctrl = h->l_ctrl.p; d.summary = h->d_summ.p;
  d.hist_cost = hist ? h->d_hist_cost.p : nullptr; d.hist_mu = hist ? h->d_hist_mu.p : nullptr;
  d.hist_accept = hist ? h->d_hist_accept.p : nullptr;
  if (h->nranks > 1) {      // what peers store into lives in the exchange arena
    d.Sg = h->peer.at(h->rank, h->peer.off_sg);
    d.pts_tri = h->peer.at(h->rank, h->peer.off_pts_tri);
  }
  return d;
}

// a8 for the large path: S block structure, this rank's landmark range and its camera-major stream
int setup_large(b2ba* h) {
  if (h->B != 1) return fail(h, B2BA_E_UNSUPPORTED, "the large (PCG) path takes one problem, not a batch");
  if (h->nranks > 1 && !h->range_set) return fail(h, B2BA_E_STATE, "nranks > 1: call b2ba_set_point_range first");
  if (!h->range_set) { h->p_lo = 0; h->p_hi = h->np_tot; }
  if (h->p_lo < 0 || h->p_hi > h->np_tot || h->p_lo > h->p_hi) return fail(h, B2BA_E_INVALID, "bad point range");
  int rc = large_build_structure(h->nc_tot, h->np_tot, h->no, h->d_pm.p, h->d_pt_ptr.p, h->h_fixed.data(), h->lix, nullptr);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "large_build_structure");
  const size_t nc = h->nc_tot, n = 6 * (size_t)h->lix.nfree, nblk = h->lix.nblk;
  if (n * sizeof(double) > 200 * 1024) return fail(h, B2BA_E_UNSUPPORTED, "PCG keeps the search direction in shared memory: at most 4266 free cameras");
  // camera-major stream, Schur pair lists and Z records cover ALL landmarks on every rank (on nranks > 1 the passes
  // over the observations are replicated work; what is sharded is S by block row and the back-substitution by landmark)
  CU(h->l_cm.alloc(h->no), "alloc"); CU(h->l_cam_ptr.alloc(nc + 1), "alloc");
  rc = large_build_cm_slice(h->nc_tot, 0, h->np_tot, h->d_pm.p, h->d_pt_ptr.p, h->l_cm.p, h->l_cam_ptr.p, nullptr);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "large_build_cm_slice");
  rc = large_build_pairs(0, h->np_tot, h->no, h->d_pm.p, h->d_pt_ptr.p, h->lix, nullptr);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "large_build_pairs");
  CU(h->l_zbuf.alloc((size_t)h->no * 24), "alloc");
  int blocks = 0;
  size_t smem = 0;
  rc = h->nranks > 1 ? large_pcg_dist_grid((int)n, h->nranks, &blocks, &smem) : large_pcg_grid((int)n, &blocks, &smem);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "large_pcg_grid");
  if (h->nranks > 1) {
    rc = large_build_rows(h->rank, h->nranks, h->lix, nullptr);
    if (rc) return cuda_fail(h, (cudaError_t)rc, "large_build_rows");
    CU(h->l_Sl.alloc((size_t)(h->lix.nlocal > 0 ? h->lix.nlocal : 1) * 36), "alloc");
    if ((rc = setup_arena(h)) != 0) return rc;
  }
  const size_t nparts = ((size_t)(h->p_hi - h->p_lo) + 127) / 128;
  size_t part = 2 * nc;
  if ((size_t)blocks * 4 > part) part = (size_t)blocks * 4;
  if (nparts * 8 > part) part = nparts * 8;
  CU(h->l_poses_tri.alloc(nc * 12), "alloc"); CU(h->l_lin_part.alloc(nc * 27 + 2), "alloc");
  CU(h->l_Dc.alloc(nc * 6), "alloc");
  CU(h->l_Sg.alloc(nblk * 36 + n), "alloc"); CU(h->l_dc.alloc(n), "alloc"); CU(h->l_r.alloc(n), "alloc");
  CU(h->l_z.alloc(n), "alloc"); CU(h->l_q.alloc(n), "alloc"); CU(h->l_Minv.alloc(6 * n), "alloc"); CU(h->l_tbuf.alloc(std::max((size_t)nblk * 6, (size_t)blocks * n)), "alloc");   // tbuf, or one n-vector per PCG CTA
  CU(h->l_part.alloc(part + 8), "alloc"); CU(h->l_sums.alloc(16), "alloc");
  if (!h->l_ctrl.p) {       // the exchange epoch lives in the control block and only grows: zero it once
    CU(h->l_ctrl.alloc(1), "alloc");
    CU(cudaMemset(h->l_ctrl.p, 0, sizeof(LargeCtrl)), "memset");
  }
  // solver choice on the large path: S block-banded (half-bandwidth <= kBandMaxBw camera blocks) -> exact banded
  // Cholesky, unless PCG was asked for explicitly; otherwise block-Jacobi PCG
  h->l_bw = -1;
  h->l_dense_ld = 0;
  if (h->opt.solver == B2BA_SOLVER_DENSE && h->lix.nfree > 0) {
    if ((int)n > large_dense_max_n()) return fail(h, B2BA_E_UNSUPPORTED, "dense Cholesky: too many free cameras for one resident CTA per 64-block");
    const int npad = large_dense_npad((int)n);
    CU(h->l_dense.alloc(large_dense_count((int)n)), "alloc"); CU(h->l_dense_winv.alloc((size_t)npad * 64), "alloc");
    CU(h->l_dense_x.alloc(npad), "alloc"); CU(h->l_dense_flags.alloc(npad / 64), "alloc");
    CU(cudaMemset(h->l_dense_flags.p, 0, (size_t)(npad / 64) * sizeof(int)), "memset");
    CU(cudaMemset(h->l_dense.p, 0, large_dense_count((int)n) * sizeof(double)), "memset");   // rhs rows 1.. stay zero
    h->l_dense_ld = npad;
  } else if (h->opt.solver != B2BA_SOLVER_PCG && h->lix.nfree > 0) {
    std::vector<int> bp(h->lix.nfree + 1), bc(h->lix.nblk);
    CU(cudaMemcpy(bp.data(), h->lix.blk_ptr.p, bp.size() * 4, cudaMemcpyDeviceToHost), "read structure");
    CU(cudaMemcpy(bc.data(), h->lix.blk_col.p, bc.size() * 4, cudaMemcpyDeviceToHost), "read structure");
    int bw = 0;
    for (int f = 0; f < h->lix.nfree; ++f)
      for (int b = bp[f]; b < bp[f + 1]; ++b) if (bc[b] - f > bw) bw = bc[b] - f;
    if (bw <= kBandMaxBw) {
      h->l_bw = bw;
      CU(h->l_band.alloc((size_t)h->lix.nfree * (bw + 1) * 36), "alloc"); CU(h->l_binv.alloc((size_t)h->lix.nfree * 6), "alloc");
      h->l_band_P = large_band_chunks(h->lix.nfree, bw);
      CU(h->l_band_ws.alloc(large_band_ws_count(h->lix.nfree, bw, h->l_band_P)), "alloc");
    }
  }
  CU(cudaMemset(h->l_lin_part.p, 0, (nc * 27 + 2) * sizeof(double)), "memset");
  CU(cudaMemset(h->l_sums.p, 0, 16 * sizeof(double)), "memset");
  CU(cudaDeviceSynchronize(), "setup_large");
  return B2BA_OK;
}

int solve_large(b2ba* h, const Params& P, bool hist, cudaStream_t s) {
  const LargeData d = make_large(h, hist);
  const bool timed = (h->opt.flags & 16) != 0;
  if (timed) {
    while (h->t_ev.size() < (size_t)P.max_iters * 7) {
      cudaEvent_t e;
      CU(cudaEventCreate(&e), "event");
      h->t_ev.push_back(e);
    }
  }
  h->t_valid = false;
  int rc = large_init_ctrl(d, P, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "large_init_ctrl");
  for (int it = 0; it < P.max_iters; ++it) {
    rc = large_iteration(d, P, h->opt.pcg_max_iters, h->opt.pcg_tol, s, timed ? h->t_ev.data() + 7 * (size_t)it : nullptr);
    if (rc) return cuda_fail(h, (cudaError_t)rc, "large_iteration");
  }
  rc = large_finish(d, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "large_finish");
  h->t_iters = P.max_iters; h->t_valid = timed;
  return B2BA_OK;        // nranks > 1: poses and landmarks are replicated state, every rank ends with all of them
}

}  // namespace

extern "C" {

int b2ba_version(void) { return 100; }

void b2ba_default_options(b2ba_options* o) {
  std::memset(o, 0, sizeof(*o));
  o->max_iters = 10; o->loss = B2BA_LOSS_TRIVIAL; o->loss_scale = 1.0; o->z_min = 1e-3;
  o->init_mu = 1e4; o->solver = B2BA_SOLVER_AUTO; o->pcg_max_iters = 200; o->pcg_tol = 1e-10;
}

int b2ba_create(b2ba_t** out, int device, const b2ba_options* opts) {
  if (!out) return B2BA_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || device < 0 || device >= ndev) return B2BA_E_CUDA;   // no CPU fallback
  DeviceGuard dg_(device);
  if (dg_.err != cudaSuccess) return B2BA_E_CUDA;
  {
    // the index builder's temporaries are stream-ordered allocations: keep them cached in the pool
    // between set_problem calls instead of returning them to the driver at every synchronisation
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      unsigned long long thr = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
  }
  b2ba* h = new b2ba();
  h->device = device;
  cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, device);
  if (opts) h->opt = *opts; else b2ba_default_options(&h->opt);
  *out = h;
  return B2BA_OK;
}

void b2ba_destroy(b2ba_t* h) {
  if (!h) return;
  DeviceGuard dg_(h->device);
  for (b2ba* k : h->kids) b2ba_destroy(k);
  for (cudaEvent_t e : h->ev_in) cudaEventDestroy(e);
  for (cudaEvent_t e : h->ev_done) cudaEventDestroy(e);
  if (h->pipe_copy) cudaStreamDestroy(h->pipe_copy);
  if (h->pipe_solve) cudaStreamDestroy(h->pipe_solve);
  if (h->pipe_solve2) cudaStreamDestroy(h->pipe_solve2);
  if (h->pinned_stage) cudaFreeHost(h->pinned_stage);
  if (h->pipe_back) cudaStreamDestroy(h->pipe_back);
  if (h->pinned_flags) cudaFreeHost(h->pinned_flags);
  for (cudaEvent_t e : h->t_ev) cudaEventDestroy(e);
  if (h->t_setup[0]) { cudaEventDestroy(h->t_setup[0]); cudaEventDestroy(h->t_setup[1]); }
  close_peers(h);
  if (h->nccl_comm && nccl_api().ok) nccl_api().CommDestroy(h->nccl_comm);
  delete h;
}

const char* b2ba_last_error(const b2ba_t* h) { return h ? h->err.c_str() : "null handle"; }

int b2ba_set_problem_batch(b2ba_t* h, int32_t B, const int32_t* cam_off, const int32_t* pt_off,
                           const int64_t* obs_off, const double* poses_cw, const double* points,
                           const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv,
                           const uint8_t* fixed_cam, const double intr[4], int mem) {
  if (!h) return B2BA_E_INVALID;
  if (B < 1 || !cam_off || !pt_off || !obs_off || !poses_cw || !points || !fixed_cam || !intr)
    return fail(h, B2BA_E_INVALID, "set_problem: null argument or B < 1");
  if (cam_off[0] != 0 || pt_off[0] != 0 || obs_off[0] != 0) return fail(h, B2BA_E_INVALID, "offsets must start at 0");
  for (int b = 0; b < B; ++b)
    if (cam_off[b + 1] < cam_off[b] || pt_off[b + 1] < pt_off[b] || obs_off[b + 1] < obs_off[b])
      return fail(h, B2BA_E_INVALID, "offsets must be non-decreasing");
  const long long no = obs_off[B];
  if (no > 0 && (!obs_cam || !obs_pt || !obs_uv)) return fail(h, B2BA_E_INVALID, "null observation arrays");
  if (no >= (1ll << 31)) return fail(h, B2BA_E_UNSUPPORTED, "more than 2^31 observations per handle");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  h->B = B; h->nc_tot = cam_off[B]; h->np_tot = pt_off[B]; h->no = no;
  h->solved = false;
  std::memcpy(h->intr, intr, 4 * sizeof(double));
  h->h_cam_off.assign(cam_off, cam_off + B + 1);
  h->h_pt_off.assign(pt_off, pt_off + B + 1);
  h->h_obs_off.assign(obs_off, obs_off + B + 1);
  const cudaMemcpyKind k = in_kind(mem);
  const size_t nc = h->nc_tot, np = h->np_tot;
  CU(h->d_cam_off.alloc(B + 1), "alloc"); CU(h->d_pt_off.alloc(B + 1), "alloc"); CU(h->d_obs_off.alloc(B + 1), "alloc");
  CU(h->d_fixed.alloc(nc), "alloc"); CU(h->d_poses.alloc(nc * 12), "alloc"); CU(h->d_poses0.alloc(nc * 12), "alloc");
  CU(h->d_pts0.alloc(np * 3), "alloc"); CU(h->d_pts1.alloc(np * 3), "alloc"); CU(h->d_ptsinit.alloc(np * 3), "alloc");
  CU(h->d_scratch.alloc(np * kPtScratch), "alloc");
  CU(h->d_obs_cam.alloc(no), "alloc"); CU(h->d_obs_pt.alloc(no), "alloc"); CU(h->d_uv.alloc(no), "alloc");
  CU(h->d_pm.alloc(no), "alloc"); CU(h->d_cm.alloc(no), "alloc");
  CU(h->d_pt_ptr.alloc(np + 1), "alloc"); CU(h->d_cam_ptr.alloc(nc + 1), "alloc"); CU(h->d_flags.alloc(2), "alloc");
  CU(h->d_sched.alloc(np), "alloc");
  CU(h->d_summ.alloc(B), "alloc");
  // asynchronous copies on the default stream (pinned caller buffers: no per-copy round trip; pageable ones are staged
  // by the runtime before the call returns); the index builder's flag read-back below waits for all of them, so no
  // caller pointer is in use after set_problem returns
  cudaStream_t cs = nullptr;
  CU(cudaMemcpyAsync(h->d_cam_off.p, cam_off, (B + 1) * sizeof(int), cudaMemcpyHostToDevice, cs), "copy offsets");
  CU(cudaMemcpyAsync(h->d_pt_off.p, pt_off, (B + 1) * sizeof(int), cudaMemcpyHostToDevice, cs), "copy offsets");
  CU(cudaMemcpyAsync(h->d_obs_off.p, obs_off, (B + 1) * sizeof(long long), cudaMemcpyHostToDevice, cs), "copy offsets");
  CU(cudaMemcpyAsync(h->d_fixed.p, fixed_cam, nc, k, cs), "copy fixed");
  CU(cudaMemcpyAsync(h->d_poses.p, poses_cw, nc * 12 * sizeof(double), k, cs), "copy poses");
  CU(cudaMemcpyAsync(h->d_pts0.p, points, np * 3 * sizeof(double), k, cs), "copy points");
  if (no) {
    CU(cudaMemcpyAsync(h->d_obs_cam.p, obs_cam, no * sizeof(int), k, cs), "copy obs_cam");
    CU(cudaMemcpyAsync(h->d_obs_pt.p, obs_pt, no * sizeof(int), k, cs), "copy obs_pt");
    CU(cudaMemcpyAsync(h->d_uv.p, obs_uv, no * sizeof(float2), k, cs), "copy obs_uv");
  }
  CU(cudaMemcpyAsync(h->d_poses0.p, h->d_poses.p, nc * 12 * sizeof(double), cudaMemcpyDeviceToDevice, cs), "keep poses");
  CU(cudaMemcpyAsync(h->d_ptsinit.p, h->d_pts0.p, np * 3 * sizeof(double), cudaMemcpyDeviceToDevice, cs), "keep points");
  // gauge / path selection needs the fixed mask on the host
  h->h_fixed.resize(nc);
  if (mem == B2BA_MEM_HOST) std::memcpy(h->h_fixed.data(), fixed_cam, nc);
  else CU(cudaMemcpy(h->h_fixed.data(), h->d_fixed.p, nc, cudaMemcpyDeviceToHost), "read fixed");
  h->max_nfree = 0; h->max_nc = 0;
  for (int b = 0; b < B; ++b) {
    int nf = 0;
    for (int c = cam_off[b]; c < cam_off[b + 1]; ++c) nf += h->h_fixed[c] ? 0 : 1;
    if (nf > h->max_nfree) h->max_nfree = nf;
    if (cam_off[b + 1] - cam_off[b] > h->max_nc) h->max_nc = cam_off[b + 1] - cam_off[b];
  }
  // a8: index.  The window kernels walk the landmarks of a window by decreasing observation count and want the
  // point-major stream in that order; the large path keeps the canonical order (by landmark id).
  const bool want_pcg = h->opt.solver == B2BA_SOLVER_PCG || h->opt.solver == B2BA_SOLVER_DENSE;   // = the large path
  h->sched_order = !want_pcg && h->max_nc <= kMaxCam && h->max_nfree <= kBigMaxFree && np > 0;
  IndexOut io{h->d_pm.p, h->d_pt_ptr.p, h->d_cm.p, h->d_cam_ptr.p, h->d_flags.p, h->sched_order ? h->d_sched.p : nullptr};
  h->t_setup_valid = false;
  if (h->opt.flags & 16) {
    if (!h->t_setup[0]) { CU(cudaEventCreate(&h->t_setup[0]), "event"); CU(cudaEventCreate(&h->t_setup[1]), "event"); }
    CU(cudaEventRecord(h->t_setup[0], nullptr), "event");
  }
  nvtxRangePushA("b2ba:index_build");
  int rc = build_index(make_index_in(h), io, nullptr);
  nvtxRangePop();
  if (rc) { cudaDeviceSynchronize(); return cuda_fail(h, (cudaError_t)rc, "build_index"); }   // copies above are done with the caller's buffers
  int flags[2] = {0, 0};
  CU(cudaMemcpy(flags, h->d_flags.p, sizeof(flags), cudaMemcpyDeviceToHost), "read index flags");
  if (flags[0]) { h->path = PATH_NONE; return fail(h, B2BA_E_INVALID, "observation index out of range"); }
  if (flags[1]) { h->path = PATH_NONE; return fail(h, B2BA_E_DUPLICATE, "duplicate (camera, point) observation"); }
  if (h->sched_order) {   // the engine keeps the landmarks in schedule order from here on
    rc = permute_pts(h, h->d_ptsinit.p, h->d_pts0.p, true, nullptr);
    if (rc) return cuda_fail(h, (cudaError_t)rc, "permute points");
    CU(cudaMemcpy(h->d_ptsinit.p, h->d_pts0.p, np * 3 * sizeof(double), cudaMemcpyDeviceToDevice), "keep points");
  }
  if (!want_pcg && h->max_nc <= kSmallMaxCam && h->max_nfree <= kSmallMaxFree) h->path = PATH_WIN_SMALL;
  else if (!want_pcg && h->max_nc <= kMaxCam && h->max_nfree <= kBigMaxFree) h->path = PATH_WIN_BIG;
  else if (h->opt.solver == B2BA_SOLVER_WINDOW)
    return fail(h, B2BA_E_UNSUPPORTED, "window solver supports at most 21 free / 24 total cameras per window");
  else h->path = PATH_LARGE;
  if (h->nranks > 1 && h->path != PATH_LARGE)
    return fail(h, B2BA_E_UNSUPPORTED, "observation sharding (nranks > 1) needs solver = B2BA_SOLVER_PCG");
  if (h->path == PATH_LARGE) {
    nvtxRangePushA("b2ba:large_structure");
    const int rc2 = setup_large(h);
    nvtxRangePop();
    if (rc2) { h->path = PATH_NONE; return rc2; }
  }
  if (h->opt.flags & 16) { CU(cudaEventRecord(h->t_setup[1], nullptr), "event"); h->t_setup_valid = true; }
  return B2BA_OK;
}

int b2ba_set_problem(b2ba_t* h, int32_t Nc, int32_t Np, int64_t No, const double* poses_cw,
                     const double* points, const int32_t* obs_cam, const int32_t* obs_pt,
                     const float* obs_uv, const uint8_t* fixed_cam, const double intr[4], int mem) {
  if (Nc < 0 || Np < 0 || No < 0) return fail(h, B2BA_E_INVALID, "negative size");
  const int32_t co[2] = {0, Nc}, po[2] = {0, Np};
  const int64_t oo[2] = {0, No};
  return b2ba_set_problem_batch(h, 1, co, po, oo, poses_cw, points, obs_cam, obs_pt, obs_uv, fixed_cam, intr, mem);
}

int b2ba_set_state(b2ba_t* h, const double* poses_cw, const double* points, int mem) {
  if (!h || h->path == PATH_NONE) return fail(h, B2BA_E_STATE, "no problem set");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  const cudaMemcpyKind k = in_kind(mem);
  if (poses_cw) {
    CU(cudaMemcpy(h->d_poses.p, poses_cw, (size_t)h->nc_tot * 12 * sizeof(double), k), "copy poses");
    CU(cudaMemcpy(h->d_poses0.p, h->d_poses.p, (size_t)h->nc_tot * 12 * sizeof(double), cudaMemcpyDeviceToDevice), "keep poses");
  }
  if (points) {
    if (h->sched_order) {
      CU(cudaMemcpy(h->d_pts1.p, points, (size_t)h->np_tot * 3 * sizeof(double), k), "copy points");
      const int rc = permute_pts(h, h->d_pts1.p, h->d_pts0.p, true, nullptr);
      if (rc) return cuda_fail(h, (cudaError_t)rc, "permute points");
    } else {
      CU(cudaMemcpy(h->d_pts0.p, points, (size_t)h->np_tot * 3 * sizeof(double), k), "copy points");
    }
    CU(cudaMemcpy(h->d_ptsinit.p, h->d_pts0.p, (size_t)h->np_tot * 3 * sizeof(double), cudaMemcpyDeviceToDevice), "keep points");
  }
  return B2BA_OK;
}

int b2ba_reset(b2ba_t* h) {
  if (!h || h->path == PATH_NONE) return fail(h, B2BA_E_STATE, "no problem set");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  cudaStream_t s = h->last_stream;
  CU(cudaMemcpyAsync(h->d_poses.p, h->d_poses0.p, (size_t)h->nc_tot * 12 * sizeof(double), cudaMemcpyDeviceToDevice, s), "reset poses");
  CU(cudaMemcpyAsync(h->d_pts0.p, h->d_ptsinit.p, (size_t)h->np_tot * 3 * sizeof(double), cudaMemcpyDeviceToDevice, s), "reset points");
  return B2BA_OK;
}

int b2ba_get_summaries(b2ba_t* h, b2ba_summary* summaries) {
  if (!h || !h->solved || !summaries) return fail(h, B2BA_E_STATE, "no solve to report");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  CU(cudaMemcpyAsync(summaries, h->d_summ.p, (size_t)h->B * sizeof(b2ba_summary), cudaMemcpyDeviceToHost, h->last_stream), "read summaries");
  CU(cudaStreamSynchronize(h->last_stream), "solve");
  if (h->path == PATH_LARGE && h->nranks > 1) {
    LargeCtrl c;
    CU(cudaMemcpy(&c, h->l_ctrl.p, sizeof(c), cudaMemcpyDeviceToHost), "read control");
    if (c.xerr) return fail(h, B2BA_E_NCCL, "peer-memory exchange timed out (a rank did not arrive): solve abandoned");
  }
  return B2BA_OK;
}

int b2ba_get_timings(b2ba_t* h, b2ba_timings* out) {
  if (!h || !out) return B2BA_E_INVALID;
  if (!h->solved || !h->t_valid) return fail(h, B2BA_E_STATE, "no timed solve (options.flags bit 4)");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  CU(cudaStreamSynchronize(h->last_stream), "solve");
  std::memset(out, 0, sizeof(*out));
  float ms = 0.f;
  if (h->t_setup_valid && cudaEventElapsedTime(&ms, h->t_setup[0], h->t_setup[1]) == cudaSuccess) out->setup_ms = ms;
  out->iterations_enqueued = h->t_iters;
  if (h->t_iters == 0) {          // window path: one kernel
    CU(cudaEventElapsedTime(&ms, h->t_ev[0], h->t_ev[1]), "elapsed");
    out->solve_ms = ms;
    return B2BA_OK;
  }
  for (int it = 0; it < h->t_iters; ++it)
    for (int p = 0; p < 6; ++p) {
      CU(cudaEventElapsedTime(&ms, h->t_ev[7 * (size_t)it + p], h->t_ev[7 * (size_t)it + p + 1]), "elapsed");
      out->phase_ms[p] += ms;
    }
  CU(cudaEventElapsedTime(&ms, h->t_ev[0], h->t_ev[7 * (size_t)(h->t_iters - 1) + 6]), "elapsed");
  out->solve_ms = ms;
  return B2BA_OK;
}

int b2ba_solve(b2ba_t* h, int32_t max_iters, b2ba_summary* summaries, void* cuda_stream) {
  if (!h || h->path == PATH_NONE) return fail(h, B2BA_E_STATE, "no problem set");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  cudaStream_t s = (cudaStream_t)cuda_stream;
  h->last_stream = s;
  const Params P = make_params(h, max_iters);
  const bool hist = (h->opt.flags & 1) != 0;
  if (hist) {
    CU(h->d_hist_cost.alloc((size_t)h->B * kHist), "alloc"); CU(h->d_hist_mu.alloc((size_t)h->B * kHist), "alloc");
    CU(h->d_hist_accept.alloc((size_t)h->B * kHist), "alloc");
    CU(cudaMemsetAsync(h->d_hist_cost.p, 0, (size_t)h->B * kHist * 8, s), "memset");
    CU(cudaMemsetAsync(h->d_hist_mu.p, 0, (size_t)h->B * kHist * 8, s), "memset");
    CU(cudaMemsetAsync(h->d_hist_accept.p, 0, (size_t)h->B * kHist * 4, s), "memset");
  }
  int rc = 0;
  if (h->path == PATH_WIN_SMALL || h->path == PATH_WIN_BIG) {
    const bool timed = (h->opt.flags & 16) != 0;
    if (timed) {
      while (h->t_ev.size() < 2) { cudaEvent_t e; CU(cudaEventCreate(&e), "event"); h->t_ev.push_back(e); }
      CU(cudaEventRecord(h->t_ev[0], s), "event");
    }
    nvtxRangePushA("b2ba:window_lm_kernel");
    rc = launch_window_path(h, make_win(h, hist, nullptr), P, s);
    nvtxRangePop();
    if (timed) CU(cudaEventRecord(h->t_ev[1], s), "event");
    h->t_iters = 0; h->t_valid = timed;
  }
  else if (h->path == PATH_LARGE) {
    if (hist) { CU(h->d_hist_cost.alloc(kHist), "alloc"); }
    rc = solve_large(h, P, hist, s);
    if (rc) return rc;
  } else return fail(h, B2BA_E_UNSUPPORTED, "path not available");
  if (rc) return cuda_fail(h, (cudaError_t)rc, "launch");
  h->solved = true;
  if (!summaries) return B2BA_OK;
  rc = b2ba_get_summaries(h, summaries);
  if (rc) return rc;
  int worst = 0;
  for (int b = 0; b < h->B; ++b) if (summaries[b].status > worst) worst = summaries[b].status;
  return worst;
}

int b2ba_get(b2ba_t* h, double* poses_cw, double* points, int mem) {
  if (!h || h->path == PATH_NONE) return fail(h, B2BA_E_STATE, "no problem set");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  cudaStream_t s = h->last_stream;
  const cudaMemcpyKind k = out_kind(mem);
  if (poses_cw) CU(cudaMemcpyAsync(poses_cw, h->d_poses.p, (size_t)h->nc_tot * 12 * sizeof(double), k, s), "read poses");
  if (points) {
    const double* src = h->d_pts0.p;
    if (h->sched_order) {   // back to the caller's landmark order (pts1 is scratch between solves)
      const int rc = permute_pts(h, h->d_pts0.p, h->d_pts1.p, false, s);
      if (rc) return cuda_fail(h, (cudaError_t)rc, "permute points");
      src = h->d_pts1.p;
    }
    CU(cudaMemcpyAsync(points, src, (size_t)h->np_tot * 3 * sizeof(double), k, s), "read points");
  }
  CU(cudaStreamSynchronize(s), "get");
  return B2BA_OK;
}

int b2ba_reproj_errors(b2ba_t* h, double* err, int mem) {
  if (!h || h->path == PATH_NONE) return fail(h, B2BA_E_STATE, "no problem set");
  if (!err && h->no > 0) return fail(h, B2BA_E_INVALID, "reproj_errors: null destination");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  cudaStream_t s = h->last_stream;
  if (h->no == 0) return B2BA_OK;
  const double* pts_lm = h->d_pts0.p;
  if (h->sched_order) {   // the observation arrays index landmarks in the caller's order
    const int rcp = permute_pts(h, h->d_pts0.p, h->d_pts1.p, false, s);
    if (rcp) return cuda_fail(h, (cudaError_t)rcp, "permute points");
    pts_lm = h->d_pts1.p;
  }
  DevBuf<double> tmp;
  double* dst = err;
  if (mem != B2BA_MEM_DEVICE) { CU(tmp.alloc((size_t)h->no), "alloc"); dst = tmp.p; }
  const int rc = obs_errors(make_index_in(h), make_params(h, 1), h->d_poses.p, pts_lm, dst, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "obs_errors");
  if (mem != B2BA_MEM_DEVICE) CU(cudaMemcpyAsync(err, tmp.p, (size_t)h->no * sizeof(double), cudaMemcpyDeviceToHost, s), "read errors");
  CU(cudaStreamSynchronize(s), "reproj_errors");
  return B2BA_OK;
}

int b2ba_cull(b2ba_t* h, double max_err, uint8_t* keep, int32_t* live, int32_t* kept_index, int64_t* n_kept,
              int apply, int mem) {
  if (!h || h->path == PATH_NONE) return fail(h, B2BA_E_STATE, "no problem set");
  if (!(max_err >= 0.0)) return fail(h, B2BA_E_INVALID, "cull: max_err must be >= 0");
  if (apply && h->nranks > 1) return fail(h, B2BA_E_UNSUPPORTED, "cull: apply is not available on a sharded (multi-rank) problem");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  cudaStream_t s = h->last_stream;
  const size_t no = (size_t)h->no, np = (size_t)h->np_tot, nc = (size_t)h->nc_tot;
  const bool dev = mem == B2BA_MEM_DEVICE;
  // the observation arrays index landmarks in the caller's order
  const double* pts_lm = h->d_pts0.p;
  if (h->sched_order) {
    const int rcp = permute_pts(h, h->d_pts0.p, h->d_pts1.p, false, s);
    if (rcp) return cuda_fail(h, (cudaError_t)rcp, "permute points");
    pts_lm = h->d_pts1.p;
  }
  DevBuf<unsigned char> t_keep, t_work;
  DevBuf<int> t_live, t_idx, t_cam, t_pt;
  DevBuf<float2> t_uv;
  DevBuf<long long> t_cnt;                       // [0] n_kept, [1 .. B] kept per window
  CU(t_cnt.alloc((size_t)h->B + 1), "alloc"); CU(t_work.alloc(cull_work_bytes(h->no)), "alloc");
  unsigned char* d_keep = nullptr;
  int *d_live = nullptr, *d_idx = nullptr;
  if (keep) { if (dev) d_keep = keep; else { CU(t_keep.alloc(no), "alloc"); d_keep = t_keep.p; } }
  if (live) { if (dev) d_live = live; else { CU(t_live.alloc(np), "alloc"); d_live = t_live.p; } }
  if (kept_index) { if (dev) d_idx = kept_index; else { CU(t_idx.alloc(no), "alloc"); d_idx = t_idx.p; } }
  if (apply) { CU(t_cam.alloc(no), "alloc"); CU(t_pt.alloc(no), "alloc"); CU(t_uv.alloc(no), "alloc"); }
  const int rc = cull_observations(make_index_in(h), make_params(h, 1), h->d_poses.p, pts_lm, max_err, d_keep, d_live,
                                   t_cnt.p + 1, apply ? t_cam.p : nullptr, apply ? t_pt.p : nullptr,
                                   apply ? t_uv.p : nullptr, d_idx, t_cnt.p, t_work.p, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "cull_kernel");
  std::vector<long long> cnt((size_t)h->B + 1);
  CU(cudaMemcpyAsync(cnt.data(), t_cnt.p, cnt.size() * sizeof(long long), cudaMemcpyDeviceToHost, s), "read counts");
  CU(cudaStreamSynchronize(s), "cull");
  if (n_kept) *n_kept = cnt[0];
  if (!dev) {
    if (keep && no) CU(cudaMemcpyAsync(keep, d_keep, no, cudaMemcpyDeviceToHost, s), "read keep");
    if (live && np) CU(cudaMemcpyAsync(live, d_live, np * sizeof(int), cudaMemcpyDeviceToHost, s), "read live");
    if (kept_index && cnt[0]) CU(cudaMemcpyAsync(kept_index, d_idx, (size_t)cnt[0] * sizeof(int), cudaMemcpyDeviceToHost, s), "read index");
    CU(cudaStreamSynchronize(s), "cull");
  }
  if (!apply) return B2BA_OK;
  // the culled problem replaces the handle's: compacted observation arrays (device), state = the present one
  DevBuf<double> t_poses, t_pts;
  DevBuf<unsigned char> t_fixed;
  CU(t_poses.alloc(nc * 12), "alloc"); CU(t_pts.alloc(np * 3), "alloc"); CU(t_fixed.alloc(nc), "alloc");
  CU(cudaMemcpyAsync(t_poses.p, h->d_poses.p, nc * 12 * sizeof(double), cudaMemcpyDeviceToDevice, s), "copy poses");
  CU(cudaMemcpyAsync(t_pts.p, pts_lm, np * 3 * sizeof(double), cudaMemcpyDeviceToDevice, s), "copy points");
  CU(cudaMemcpyAsync(t_fixed.p, h->d_fixed.p, nc, cudaMemcpyDeviceToDevice, s), "copy fixed");
  CU(cudaStreamSynchronize(s), "cull");
  const std::vector<int> cam_off = h->h_cam_off, pt_off = h->h_pt_off;
  std::vector<long long> obs_off((size_t)h->B + 1, 0);
  for (int b = 0; b < h->B; ++b) obs_off[b + 1] = obs_off[b] + cnt[1 + b];
  double intr[4];
  std::memcpy(intr, h->intr, sizeof(intr));
  static_assert(sizeof(long long) == sizeof(int64_t), "obs_off");
  return b2ba_set_problem_batch(h, h->B, cam_off.data(), pt_off.data(), reinterpret_cast<const int64_t*>(obs_off.data()),
                                t_poses.p, t_pts.p, t_cam.p, t_pt.p, reinterpret_cast<const float*>(t_uv.p), t_fixed.p,
                                intr, B2BA_MEM_DEVICE);
}

int b2ba_triangulate(b2ba_t* h, int32_t N, const double P1[12], const double P2[12], const float* uv1,
                     const float* uv2, double min_depth, double max_reproj_err, double* X, uint8_t* keep, int mem) {
  if (!h) return B2BA_E_INVALID;
  if (N < 0 || !P1 || !P2 || (N > 0 && (!uv1 || !uv2 || !X || !keep))) return fail(h, B2BA_E_INVALID, "triangulate: null argument");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  if (N == 0) return B2BA_OK;
  cudaStream_t s = nullptr;
  if (mem == B2BA_MEM_DEVICE) {
    const int rc = launch_triangulate(N, P1, P2, (const float2*)uv1, (const float2*)uv2, min_depth, max_reproj_err, X, keep, s);
    if (rc) return cuda_fail(h, (cudaError_t)rc, "triangulate_kernel");
    CU(cudaStreamSynchronize(s), "triangulate");
    return B2BA_OK;
  }
  DevBuf<float2> a, b;
  DevBuf<double> x;
  DevBuf<unsigned char> k;
  CU(a.alloc(N), "alloc"); CU(b.alloc(N), "alloc"); CU(x.alloc((size_t)N * 3), "alloc"); CU(k.alloc(N), "alloc");
  CU(cudaMemcpyAsync(a.p, uv1, (size_t)N * sizeof(float2), cudaMemcpyHostToDevice, s), "copy uv1");
  CU(cudaMemcpyAsync(b.p, uv2, (size_t)N * sizeof(float2), cudaMemcpyHostToDevice, s), "copy uv2");
  const int rc = launch_triangulate(N, P1, P2, a.p, b.p, min_depth, max_reproj_err, x.p, k.p, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "triangulate_kernel");
  CU(cudaMemcpyAsync(X, x.p, (size_t)N * 3 * sizeof(double), cudaMemcpyDeviceToHost, s), "read X");
  CU(cudaMemcpyAsync(keep, k.p, (size_t)N, cudaMemcpyDeviceToHost, s), "read keep");
  CU(cudaStreamSynchronize(s), "triangulate");
  return B2BA_OK;
}

int b2ba_refine_poses(b2ba_t* h, int32_t B, const int64_t* obs_off, double* poses_cw, const double* points,
                      const float* obs_uv, const double intr[4], int32_t max_iters, b2ba_summary* summaries, int mem) {
  if (!h) return B2BA_E_INVALID;
  if (B < 0 || !obs_off || (B > 0 && (!poses_cw || !intr))) return fail(h, B2BA_E_INVALID, "refine_poses: null argument");
  for (int b = 0; b < B; ++b)
    if (obs_off[b + 1] < obs_off[b] || obs_off[0] != 0) return fail(h, B2BA_E_INVALID, "refine_poses: obs_off must start at 0 and be non-decreasing");
  const int64_t no = B > 0 ? obs_off[B] : 0;
  if (no > 0 && (!points || !obs_uv)) return fail(h, B2BA_E_INVALID, "refine_poses: null argument");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  h->r_nb = B;
  if (B == 0) return B2BA_OK;
  const cudaMemcpyKind ki = in_kind(mem), ko = out_kind(mem);
  cudaStream_t s = nullptr;
  CU(h->r_off.alloc(B + 1), "alloc"); CU(h->r_poses.alloc((size_t)B * 12), "alloc"); CU(h->r_summ.alloc(B), "alloc");
  CU(h->r_pts.alloc((size_t)no * 3), "alloc"); CU(h->r_uv.alloc(no), "alloc");
  static_assert(sizeof(long long) == sizeof(int64_t), "offset type");
  CU(cudaMemcpyAsync(h->r_off.p, obs_off, (size_t)(B + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s), "copy offsets");
  CU(cudaMemcpyAsync(h->r_poses.p, poses_cw, (size_t)B * 12 * sizeof(double), ki, s), "copy poses");
  if (no) {
    CU(cudaMemcpyAsync(h->r_pts.p, points, (size_t)no * 3 * sizeof(double), ki, s), "copy points");
    CU(cudaMemcpyAsync(h->r_uv.p, obs_uv, (size_t)no * sizeof(float2), ki, s), "copy uv");
  }
  PoseData d{};
  d.nb = B; d.obs_off = h->r_off.p; d.poses = h->r_poses.p; d.points = h->r_pts.p; d.obs_uv = h->r_uv.p;
  d.summaries = h->r_summ.p;
  if (h->opt.flags & 1) {
    CU(h->r_hist_cost.alloc((size_t)B * kPoseHist), "alloc"); CU(h->r_hist_mu.alloc((size_t)B * kPoseHist), "alloc");
    CU(h->r_hist_accept.alloc((size_t)B * kPoseHist), "alloc");
    CU(cudaMemsetAsync(h->r_hist_cost.p, 0, (size_t)B * kPoseHist * 8, s), "memset");
    CU(cudaMemsetAsync(h->r_hist_mu.p, 0, (size_t)B * kPoseHist * 8, s), "memset");
    CU(cudaMemsetAsync(h->r_hist_accept.p, 0, (size_t)B * kPoseHist * 4, s), "memset");
    d.hist_cost = h->r_hist_cost.p; d.hist_mu = h->r_hist_mu.p; d.hist_accept = h->r_hist_accept.p;
  }
  Params P = make_params(h, max_iters);
  P.fx = intr[0]; P.fy = intr[1]; P.cx = intr[2]; P.cy = intr[3];
  const int rc = launch_pose_lm(d, P, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "pose_lm_kernel");
  CU(cudaMemcpyAsync(poses_cw, h->r_poses.p, (size_t)B * 12 * sizeof(double), ko, s), "read poses");
  std::vector<b2ba_summary> sm(B);
  CU(cudaMemcpyAsync(sm.data(), h->r_summ.p, (size_t)B * sizeof(b2ba_summary), cudaMemcpyDeviceToHost, s), "read summaries");
  CU(cudaStreamSynchronize(s), "refine_poses");
  int worst = 0;
  for (int b = 0; b < B; ++b) {
    if (summaries) summaries[b] = sm[b];
    if (sm[b].status > worst) worst = sm[b].status;
  }
  return worst;
}

int b2ba_track_pose(b2ba_t* h, int64_t N, double* pose_cw, const double* points, const float* obs_uv,
                    const double intr[4], double reproj_err, int32_t rounds, int32_t iters_per_round,
                    uint8_t* inliers, int32_t* n_inliers, b2ba_summary* summary, int mem) {
  if (!h) return B2BA_E_INVALID;
  if (N < 0 || !pose_cw || !intr || !n_inliers || (N > 0 && (!points || !obs_uv || !inliers)))
    return fail(h, B2BA_E_INVALID, "track_pose: null argument");
  if (!(reproj_err > 0.0) || rounds < 1) return fail(h, B2BA_E_INVALID, "track_pose: reproj_err > 0 and rounds >= 1 required");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  const cudaMemcpyKind ki = in_kind(mem), ko = out_kind(mem);
  cudaStream_t s = nullptr;
  const size_t no = (size_t)N;
  const long long off[2] = {0, (long long)N};
  CU(h->r_off.alloc(2), "alloc"); CU(h->r_poses.alloc(12), "alloc"); CU(h->r_summ.alloc(1), "alloc");
  CU(h->r_pts.alloc(no * 3), "alloc"); CU(h->r_uv.alloc(no), "alloc"); CU(h->r_mask.alloc(no), "alloc"); CU(h->r_cnt.alloc(1), "alloc");
  CU(cudaMemcpyAsync(h->r_off.p, off, sizeof(off), cudaMemcpyHostToDevice, s), "copy offsets");
  CU(cudaMemcpyAsync(h->r_poses.p, pose_cw, 12 * sizeof(double), ki, s), "copy pose");
  if (no) {
    CU(cudaMemcpyAsync(h->r_pts.p, points, no * 3 * sizeof(double), ki, s), "copy points");
    CU(cudaMemcpyAsync(h->r_uv.p, obs_uv, no * sizeof(float2), ki, s), "copy uv");
    CU(cudaMemsetAsync(h->r_mask.p, 1, no, s), "memset");
  }
  CU(cudaMemsetAsync(h->r_cnt.p, 0, sizeof(int), s), "memset");
  PoseData d{};
  d.nb = 1; d.obs_off = h->r_off.p; d.poses = h->r_poses.p; d.points = h->r_pts.p; d.obs_uv = h->r_uv.p;
  d.summaries = h->r_summ.p;
  Params P = make_params(h, iters_per_round);
  P.fx = intr[0]; P.fy = intr[1]; P.cx = intr[2]; P.cy = intr[3];
  const int rc = launch_pose_track(d, P, reproj_err, rounds, h->r_mask.p, h->r_cnt.p, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "pose_track_kernel");
  CU(cudaMemcpyAsync(pose_cw, h->r_poses.p, 12 * sizeof(double), ko, s), "read pose");
  if (no) CU(cudaMemcpyAsync(inliers, h->r_mask.p, no, ko, s), "read inliers");
  b2ba_summary sm;
  int cnt = 0;
  CU(cudaMemcpyAsync(&sm, h->r_summ.p, sizeof(sm), cudaMemcpyDeviceToHost, s), "read summary");
  CU(cudaMemcpyAsync(&cnt, h->r_cnt.p, sizeof(int), cudaMemcpyDeviceToHost, s), "read count");
  CU(cudaStreamSynchronize(s), "track_pose");
  *n_inliers = cnt;
  if (summary) *summary = sm;
  return sm.status;
}

int b2ba_ransac_pose(b2ba_t* h, int64_t N, double* pose_cw, const double* points, const float* obs_uv,
                     const double intr[4], double reproj_err, int32_t n_hyp, int32_t hyp_iters, uint64_t seed,
                     int32_t refine_iters, uint8_t* inliers, int32_t* n_inliers, int32_t* best_consensus,
                     b2ba_summary* summary, int mem) {
  if (!h) return B2BA_E_INVALID;
  if (N < 0 || !pose_cw || !intr || !n_inliers || (N > 0 && (!points || !obs_uv || !inliers)))
    return fail(h, B2BA_E_INVALID, "ransac_pose: null argument");
  if (!(reproj_err > 0.0) || n_hyp < 1 || hyp_iters < 1) return fail(h, B2BA_E_INVALID, "ransac_pose: reproj_err > 0, n_hyp >= 1, hyp_iters >= 1 required");
  *n_inliers = 0;
  if (best_consensus) *best_consensus = 0;
  if (N < 6) {                                   // nothing to vote on: the caller treats it as a failure
    if (summary) { std::memset(summary, 0, sizeof(*summary)); summary->status = B2BA_NUMERICAL_FAILURE; }
    if (N > 0 && mem == B2BA_MEM_HOST) std::memset(inliers, 0, (size_t)N);
    return B2BA_NUMERICAL_FAILURE;
  }
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  const cudaMemcpyKind ki = in_kind(mem), ko = out_kind(mem);
  cudaStream_t s = nullptr;
  const size_t no = (size_t)N, nh = (size_t)n_hyp;
  const long long off[2] = {0, (long long)N};
  CU(h->r_off.alloc(2), "alloc"); CU(h->r_poses.alloc(12), "alloc"); CU(h->r_summ.alloc(1), "alloc");
  CU(h->r_pts.alloc(no * 3), "alloc"); CU(h->r_uv.alloc(no), "alloc"); CU(h->r_mask.alloc(no), "alloc"); CU(h->r_cnt.alloc(1), "alloc");
  CU(h->q_pts.alloc(nh * kRansacSample * 3), "alloc"); CU(h->q_uv.alloc(nh * kRansacSample), "alloc");
  CU(h->q_poses.alloc(nh * 12), "alloc"); CU(h->q_off.alloc(nh + 1), "alloc"); CU(h->q_idx.alloc(nh * kRansacSample), "alloc");
  CU(h->q_summ.alloc(nh), "alloc"); CU(h->q_score.alloc(nh), "alloc"); CU(h->q_best.alloc(2), "alloc");
  CU(cudaMemcpyAsync(h->r_off.p, off, sizeof(off), cudaMemcpyHostToDevice, s), "copy offsets");
  CU(cudaMemcpyAsync(h->r_poses.p, pose_cw, 12 * sizeof(double), ki, s), "copy pose");
  CU(cudaMemcpyAsync(h->r_pts.p, points, no * 3 * sizeof(double), ki, s), "copy points");
  CU(cudaMemcpyAsync(h->r_uv.p, obs_uv, no * sizeof(float2), ki, s), "copy uv");
  CU(cudaMemsetAsync(h->r_cnt.p, 0, sizeof(int), s), "memset");
  PoseData d{};
  d.nb = 1; d.obs_off = h->r_off.p; d.poses = h->r_poses.p; d.points = h->r_pts.p; d.obs_uv = h->r_uv.p;
  d.summaries = h->r_summ.p;
  Params P = make_params(h, refine_iters);
  P.fx = intr[0]; P.fy = intr[1]; P.cx = intr[2]; P.cy = intr[3];
  RansacData r{n_hyp, hyp_iters, (long long)N, (unsigned long long)seed, h->q_pts.p, h->q_uv.p, h->q_poses.p, h->q_off.p,
               h->q_idx.p, h->q_summ.p, h->q_score.p, h->q_best.p};
  int rc = launch_pose_ransac(r, d, P, reproj_err, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "pose ransac kernels");
  rc = launch_pose_track(d, P, reproj_err, 2, h->r_mask.p, h->r_cnt.p, s, 1);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "pose_track_kernel");
  CU(cudaMemcpyAsync(pose_cw, h->r_poses.p, 12 * sizeof(double), ko, s), "read pose");
  CU(cudaMemcpyAsync(inliers, h->r_mask.p, no, ko, s), "read inliers");
  b2ba_summary sm;
  int cnt = 0, best[2] = {0, 0};
  CU(cudaMemcpyAsync(&sm, h->r_summ.p, sizeof(sm), cudaMemcpyDeviceToHost, s), "read summary");
  CU(cudaMemcpyAsync(&cnt, h->r_cnt.p, sizeof(int), cudaMemcpyDeviceToHost, s), "read count");
  CU(cudaMemcpyAsync(best, h->q_best.p, sizeof(best), cudaMemcpyDeviceToHost, s), "read best");
  CU(cudaStreamSynchronize(s), "ransac_pose");
  *n_inliers = cnt;
  if (best_consensus) *best_consensus = best[1];
  if (summary) *summary = sm;
  return sm.status;
}

// Pipelined batch solve (C4 end to end): the windows are cut into `chunks` contiguous chunks; chunk i's host->device
// copies (copy stream), index build + LM solve (solve stream) and device->host read-back (read-back stream)
// overlap with the neighbouring chunks', so the copies hide behind the solves.  Host buffers only; pinned
// memory makes the copies truly asynchronous.  Every chunk is an independent sub-problem (own index arrays).
int b2ba_optimize_batch(b2ba_t* h, int32_t B, const int32_t* cam_off, const int32_t* pt_off, const int64_t* obs_off,
                        const double* poses_cw, const double* points, const int32_t* obs_cam, const int32_t* obs_pt,
                        const float* obs_uv, const uint8_t* fixed_cam, const double intr[4], int32_t max_iters,
                        double* out_poses_cw, double* out_points, b2ba_summary* summaries, int32_t chunks) {
  if (!h) return B2BA_E_INVALID;
  if (B < 1 || !cam_off || !pt_off || !obs_off || !poses_cw || !points || !fixed_cam || !intr || !out_poses_cw ||
      !out_points || !summaries)
    return fail(h, B2BA_E_INVALID, "optimize_batch: null argument or B < 1");
  if (h->nranks > 1) return fail(h, B2BA_E_UNSUPPORTED, "optimize_batch is single-GPU (windows shard across ranks on the host side)");
  if (cam_off[0] != 0 || pt_off[0] != 0 || obs_off[0] != 0) return fail(h, B2BA_E_INVALID, "offsets must start at 0");
  for (int b = 0; b < B; ++b)
    if (cam_off[b + 1] < cam_off[b] || pt_off[b + 1] < pt_off[b] || obs_off[b + 1] < obs_off[b])
      return fail(h, B2BA_E_INVALID, "offsets must be non-decreasing");
  if (obs_off[B] > 0 && (!obs_cam || !obs_pt || !obs_uv)) return fail(h, B2BA_E_INVALID, "null observation arrays");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  if (!h->pipe_copy) {
    CU(cudaStreamCreateWithFlags(&h->pipe_copy, cudaStreamNonBlocking), "stream");
    CU(cudaStreamCreateWithFlags(&h->pipe_solve, cudaStreamNonBlocking), "stream");
    CU(cudaStreamCreateWithFlags(&h->pipe_solve2, cudaStreamNonBlocking), "stream");
    CU(cudaStreamCreateWithFlags(&h->pipe_back, cudaStreamNonBlocking), "stream");
    int regs = 0, smem = 0, per_sm = 0, sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
    if (window_kernel_attrs(0, &regs, &smem, &per_sm) == 0) h->wave_small = per_sm * sms;
    if (window_kernel_attrs(1, &regs, &smem, &per_sm) == 0) h->wave_big = per_sm * sms;
  }
  // path of the whole batch (every chunk runs the same kernel family)
  int max_nc = 0, max_nfree = 0;
  for (int b = 0; b < B; ++b) {
    int nf = 0;
    for (int q = cam_off[b]; q < cam_off[b + 1]; ++q) nf += fixed_cam[q] ? 0 : 1;
    if (nf > max_nfree) max_nfree = nf;
    if (cam_off[b + 1] - cam_off[b] > max_nc) max_nc = cam_off[b + 1] - cam_off[b];
  }
  Path path;
  if (max_nc <= kSmallMaxCam && max_nfree <= kSmallMaxFree) path = PATH_WIN_SMALL;
  else if (max_nc <= kMaxCam && max_nfree <= kBigMaxFree) path = PATH_WIN_BIG;
  else return fail(h, B2BA_E_UNSUPPORTED, "optimize_batch takes window problems (<= 24 cameras, <= 21 free per window)");
  // chunk boundaries: with chunks <= 0 they fall on multiples of half a launch wave: a first chunk of half a wave (its
  // upload is the only one the solve cannot hide), then one wave per chunk (the last, short chunk keeps the final
  // read-back short); few windows -> one chunk on the cluster (latency) path
  std::vector<int> cut{0};
  const int wave = path == PATH_WIN_SMALL ? h->wave_small : h->wave_big;
  if (B <= 18) cut.push_back(B);
  else if (chunks > 0) { const int K = chunks > B ? B : chunks; for (int c = 1; c <= K; ++c) cut.push_back((int)((long long)B * c / K)); }
  else if (wave <= 0 || B <= 2 * wave) cut.push_back(B);
  else {
    // B2BA_CHUNK_PLAN="first,step" in quarter waves (default 2,4: half a wave, then one wave per chunk; measured with
    // tools/e2e_chunks.py: 65.7 ms against 67.1 ms for 4,8 on the C4 batch)
    int q_first = 2, q_step = 4;
    if (const char* e = getenv("B2BA_CHUNK_PLAN")) { if (sscanf(e, "%d,%d", &q_first, &q_step) != 2 || q_first < 1 || q_step < 1) { q_first = 2; q_step = 4; } }
    int w = wave * q_first / 4;
    while (w < B) { cut.push_back(w); w += wave * q_step / 4; }
    cut.push_back(B);
  }
  const int K = (int)cut.size() - 1;
  while ((int)h->kids.size() < K) {
    b2ba* k = nullptr;
    const int rc = b2ba_create(&k, h->device, &h->opt);
    if (rc) return fail(h, rc, "optimize_batch: child engine");
    h->kids.push_back(k);
    cudaEvent_t e0, e1;
    CU(cudaEventCreateWithFlags(&e0, cudaEventDisableTiming), "event");
    CU(cudaEventCreateWithFlags(&e1, cudaEventDisableTiming), "event");
    h->ev_in.push_back(e0); h->ev_done.push_back(e1);
    if (h->pinned_flags) { cudaFreeHost(h->pinned_flags); h->pinned_flags = nullptr; }
  }
  if (!h->pinned_flags) CU(cudaMallocHost((void**)&h->pinned_flags, sizeof(int) * 2 * h->kids.size()), "pinned flags");
  const size_t off_bytes = ((size_t)B + K) * (2 * sizeof(int) + sizeof(long long));
  const size_t need = (size_t)B * sizeof(b2ba_summary) + off_bytes + 64;
  if (need > h->pinned_stage_bytes) {
    if (h->pinned_stage) cudaFreeHost(h->pinned_stage);
    h->pinned_stage = nullptr; h->pinned_stage_bytes = 0;
    CU(cudaMallocHost((void**)&h->pinned_stage, need), "pinned staging");
    h->pinned_stage_bytes = need;
  }
  b2ba_summary* st_summ = reinterpret_cast<b2ba_summary*>(h->pinned_stage);
  long long* st_obs = reinterpret_cast<long long*>(h->pinned_stage + (((size_t)B * sizeof(b2ba_summary) + 15) / 16) * 16);
  int* st_cam = reinterpret_cast<int*>(st_obs + B + K);
  int* st_pt = st_cam + B + K;
  const Params P = make_params(h, max_iters);
  h->path = PATH_NONE;                       // results live in the caller's buffers, not in this handle
  h->solved = false;
  // From the first enqueue on, copies and kernels of earlier chunks read the caller's inputs and write the caller's
  // outputs asynchronously: EVERY exit -- errors included -- first waits for all four streams, so that no work of this
  // call touches a caller buffer after it has returned.
  // B2BA_E2E_TRACE=1 (developer knob): timing events per chunk, printed to stderr after the call
  static const bool trace = [] { const char* e = getenv("B2BA_E2E_TRACE"); return e && atoi(e) != 0; }();
  std::vector<cudaEvent_t> tev;
  auto mark = [&](cudaStream_t st) { if (!trace) return; cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); tev.push_back(e); };
  mark(h->pipe_copy);
  auto bail = [&](int rc) {
    cudaStreamSynchronize(h->pipe_copy); cudaStreamSynchronize(h->pipe_solve);
    cudaStreamSynchronize(h->pipe_solve2); cudaStreamSynchronize(h->pipe_back);
    for (cudaEvent_t e : tev) cudaEventDestroy(e);
    tev.clear();
    return rc;
  };
#define CUB(x, where) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return bail(cuda_fail(h, e_, where)); } while (0)
  size_t stpos = 0;
  for (int c = 0; c < K; ++c) {
    b2ba* k = h->kids[c];
    k->opt = h->opt;
    if (B > 18) k->opt.flags |= 2;           // the batch decides the kernel family, not the chunk size
    const int w0 = cut[c], w1 = cut[c + 1], nb = w1 - w0;
    const size_t c0 = cam_off[w0], nc = cam_off[w1] - c0, p0 = pt_off[w0], np = pt_off[w1] - p0;
    const long long o0 = obs_off[w0], no = obs_off[w1] - o0;
    if (no >= (1ll << 31)) return bail(fail(h, B2BA_E_UNSUPPORTED, "more than 2^31 observations per chunk"));
    k->B = nb; k->nc_tot = (int)nc; k->np_tot = (int)np; k->no = no; k->path = path;
    std::memcpy(k->intr, intr, 4 * sizeof(double));
    int* sc = st_cam + stpos; int* sp = st_pt + stpos; long long* so = st_obs + stpos;
    stpos += nb + 1;
    for (int b = 0; b <= nb; ++b) {
      sc[b] = cam_off[w0 + b] - (int)c0; sp[b] = pt_off[w0 + b] - (int)p0; so[b] = obs_off[w0 + b] - o0;
    }
    CUB(k->d_cam_off.alloc(nb + 1), "alloc"); CUB(k->d_pt_off.alloc(nb + 1), "alloc"); CUB(k->d_obs_off.alloc(nb + 1), "alloc");
    CUB(k->d_fixed.alloc(nc), "alloc"); CUB(k->d_poses.alloc(nc * 12), "alloc");
    CUB(k->d_pts0.alloc(np * 3), "alloc"); CUB(k->d_pts1.alloc(np * 3), "alloc"); CUB(k->d_scratch.alloc(np * kPtScratch), "alloc");
    CUB(k->d_obs_cam.alloc(no), "alloc"); CUB(k->d_obs_pt.alloc(no), "alloc"); CUB(k->d_uv.alloc(no), "alloc");
    CUB(k->d_pm.alloc(no), "alloc"); CUB(k->d_cm.alloc(no), "alloc");
    CUB(k->d_pt_ptr.alloc(np + 1), "alloc"); CUB(k->d_cam_ptr.alloc(nc + 1), "alloc"); CUB(k->d_flags.alloc(2), "alloc");
    CUB(k->d_sched.alloc(np), "alloc");
    CUB(k->d_summ.alloc(nb), "alloc");
    cudaStream_t cs = h->pipe_copy, ks = (c & 1) ? h->pipe_solve2 : h->pipe_solve, bs = h->pipe_back;
    const cudaMemcpyKind hd = cudaMemcpyHostToDevice, dh = cudaMemcpyDeviceToHost;
    // nothing of an earlier call is in flight: every call ends with a synchronisation of all streams
    CUB(cudaMemcpyAsync(k->d_cam_off.p, sc, (nb + 1) * sizeof(int), hd, cs), "copy");
    CUB(cudaMemcpyAsync(k->d_pt_off.p, sp, (nb + 1) * sizeof(int), hd, cs), "copy");
    CUB(cudaMemcpyAsync(k->d_obs_off.p, so, (nb + 1) * sizeof(long long), hd, cs), "copy");
    CUB(cudaMemcpyAsync(k->d_fixed.p, fixed_cam + c0, nc, hd, cs), "copy");
    CUB(cudaMemcpyAsync(k->d_poses.p, poses_cw + c0 * 12, nc * 12 * sizeof(double), hd, cs), "copy");
    k->sched_order = np > 0;   // landmarks arrive in the caller's order (pts1) and are put into schedule order (pts0)
    CUB(cudaMemcpyAsync(k->sched_order ? k->d_pts1.p : k->d_pts0.p, points + p0 * 3, np * 3 * sizeof(double), hd, cs), "copy");
    if (no) {
      CUB(cudaMemcpyAsync(k->d_obs_cam.p, obs_cam + o0, no * sizeof(int), hd, cs), "copy");
      CUB(cudaMemcpyAsync(k->d_obs_pt.p, obs_pt + o0, no * sizeof(int), hd, cs), "copy");
      CUB(cudaMemcpyAsync(k->d_uv.p, obs_uv + 2 * o0, no * sizeof(float2), hd, cs), "copy");
    }
    CUB(cudaEventRecord(h->ev_in[c], cs), "event");
    mark(cs);
    CUB(cudaStreamWaitEvent(ks, h->ev_in[c], 0), "wait");
    mark(ks);
    IndexOut io{k->d_pm.p, k->d_pt_ptr.p, k->d_cm.p, k->d_cam_ptr.p, k->d_flags.p, k->sched_order ? k->d_sched.p : nullptr};
    IndexIn kin = make_index_in(k);
    for (int b = 0; b < nb; ++b) { kin.max_np = std::max(kin.max_np, sp[b + 1] - sp[b]); kin.max_nc = std::max(kin.max_nc, sc[b + 1] - sc[b]); }
    int rc = build_index(kin, io, ks);
    if (rc) return bail(cuda_fail(h, (cudaError_t)rc, "build_index"));
    if (k->sched_order && (rc = permute_pts(k, k->d_pts1.p, k->d_pts0.p, true, ks)) != 0) return bail(cuda_fail(h, (cudaError_t)rc, "permute"));
    mark(ks);
    rc = launch_window_path(k, make_win(k, false, nullptr), P, ks);
    if (rc) return bail(cuda_fail(h, (cudaError_t)rc, "launch"));
    if (k->sched_order && (rc = permute_pts(k, k->d_pts0.p, k->d_pts1.p, false, ks)) != 0) return bail(cuda_fail(h, (cudaError_t)rc, "permute"));
    CUB(cudaEventRecord(h->ev_done[c], ks), "event");
    mark(ks);
    CUB(cudaStreamWaitEvent(bs, h->ev_done[c], 0), "wait");
    CUB(cudaMemcpyAsync(out_poses_cw + c0 * 12, k->d_poses.p, nc * 12 * sizeof(double), dh, bs), "read");
    CUB(cudaMemcpyAsync(out_points + p0 * 3, k->sched_order ? k->d_pts1.p : k->d_pts0.p, np * 3 * sizeof(double), dh, bs), "read");
    CUB(cudaMemcpyAsync(st_summ + w0, k->d_summ.p, nb * sizeof(b2ba_summary), dh, bs), "read");
    CUB(cudaMemcpyAsync(h->pinned_flags + 2 * c, k->d_flags.p, 2 * sizeof(int), dh, bs), "read");
    mark(bs);
  }
#undef CUB
  {
    const cudaError_t e0 = cudaStreamSynchronize(h->pipe_copy), e1 = cudaStreamSynchronize(h->pipe_solve),
                      e2 = cudaStreamSynchronize(h->pipe_solve2), e3 = cudaStreamSynchronize(h->pipe_back);
    const cudaError_t ee = e0 != cudaSuccess ? e0 : e1 != cudaSuccess ? e1 : e2 != cudaSuccess ? e2 : e3;
    if (ee != cudaSuccess) return cuda_fail(h, ee, "optimize_batch");
  }
  if (trace && tev.size() == (size_t)(1 + 5 * K)) {
    for (int c = 0; c < K; ++c) {
      float t[5];
      for (int j = 0; j < 5; ++j) cudaEventElapsedTime(&t[j], tev[0], tev[1 + 5 * c + j]);
      fprintf(stderr, "chunk %d (%d windows): copied %.2f  solve stream free %.2f  index+permute done %.2f  solved %.2f  read back %.2f ms\n",
              c, cut[c + 1] - cut[c], t[0], t[1], t[2], t[3], t[4]);
    }
  }
  for (cudaEvent_t e : tev) cudaEventDestroy(e);
  for (int c = 0; c < K; ++c) {
    if (h->pinned_flags[2 * c]) return fail(h, B2BA_E_INVALID, "observation index out of range");
    if (h->pinned_flags[2 * c + 1]) return fail(h, B2BA_E_DUPLICATE, "duplicate (camera, point) observation");
  }
  std::memcpy(summaries, st_summ, (size_t)B * sizeof(b2ba_summary));
  int worst = 0;
  for (int b = 0; b < B; ++b) if (summaries[b].status > worst) worst = summaries[b].status;
  return worst;
}

int b2ba_create_comm(b2ba_t* h, const void* nccl_unique_id, int rank, int nranks) {
  if (!h || !nccl_unique_id || nranks < 1 || rank < 0 || rank >= nranks) return fail(h, B2BA_E_INVALID, "create_comm: bad arguments");
  NcclApi& api = nccl_api();
  if (!api.ok) return fail(h, B2BA_E_NCCL, "libnccl.so.2 could not be loaded");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  if (h->nccl_comm) { api.CommDestroy(h->nccl_comm); h->nccl_comm = nullptr; }
  NcclApi::UniqueId id;
  std::memcpy(&id, nccl_unique_id, sizeof(id));
  const int rc = api.CommInitRank(&h->nccl_comm, nranks, id, rank);
  if (rc != 0) return fail(h, B2BA_E_NCCL, std::string("ncclCommInitRank: ") + (api.GetErrorString ? api.GetErrorString(rc) : "error"));
  h->rank = rank; h->nranks = nranks;
  return B2BA_OK;
}

int b2ba_nccl_unique_id(void* out128) {
  NcclApi& api = nccl_api();
  if (!api.ok || !out128) return B2BA_E_NCCL;
  NcclApi::UniqueId id;
  if (api.GetUniqueId(&id) != 0) return B2BA_E_NCCL;
  std::memcpy(out128, &id, sizeof(id));
  return B2BA_OK;
}

int b2ba_set_point_range(b2ba_t* h, int32_t p_lo, int32_t p_hi) {
  if (!h || p_lo < 0 || p_hi < p_lo) return fail(h, B2BA_E_INVALID, "set_point_range: bad range");
  h->p_lo = p_lo; h->p_hi = p_hi; h->range_set = true;
  return B2BA_OK;
}

namespace {

// one LM iteration at init_mu from the current state with the stage dump attached; state restored
int run_stage(b2ba* h, std::vector<double>& out, WinDbg& L) {
  if (h->B != 1 || (h->path != PATH_WIN_SMALL && h->path != PATH_WIN_BIG))
    return fail(h, B2BA_E_UNSUPPORTED, "stage dumps need a single-window problem on the window path");
  int nfree = 0;
  for (int c = 0; c < h->nc_tot; ++c) nfree += h->h_fixed[c] ? 0 : 1;
  L = WinDbg{h->nc_tot, h->np_tot, 6 * nfree};
  CU(h->d_dbg.alloc(L.total()), "alloc dbg");
  CU(cudaMemset(h->d_dbg.p, 0, L.total() * sizeof(double)), "memset dbg");
  DevBuf<double> sp, sx;
  DevBuf<b2ba_summary> ss;
  CU(sp.alloc((size_t)h->nc_tot * 12), "alloc"); CU(sx.alloc((size_t)h->np_tot * 3), "alloc"); CU(ss.alloc(1), "alloc");
  CU(cudaMemcpy(sp.p, h->d_poses.p, (size_t)h->nc_tot * 12 * 8, cudaMemcpyDeviceToDevice), "save");
  CU(cudaMemcpy(sx.p, h->d_pts0.p, (size_t)h->np_tot * 3 * 8, cudaMemcpyDeviceToDevice), "save");
  CU(cudaMemcpy(ss.p, h->d_summ.p, sizeof(b2ba_summary), cudaMemcpyDeviceToDevice), "save");
  const Params P = make_params(h, 1);
  const WinData d = make_win(h, false, h->d_dbg.p);
  int rc = h->path == PATH_WIN_SMALL ? launch_window_small(d, P, nullptr) : launch_window_big(d, P, nullptr);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "stage launch");
  CU(cudaDeviceSynchronize(), "stage kernel");
  out.resize(L.total());
  CU(cudaMemcpy(out.data(), h->d_dbg.p, L.total() * 8, cudaMemcpyDeviceToHost), "read dbg");
  CU(cudaMemcpy(h->d_poses.p, sp.p, (size_t)h->nc_tot * 12 * 8, cudaMemcpyDeviceToDevice), "restore");
  CU(cudaMemcpy(h->d_pts0.p, sx.p, (size_t)h->np_tot * 3 * 8, cudaMemcpyDeviceToDevice), "restore");
  CU(cudaMemcpy(h->d_summ.p, ss.p, sizeof(b2ba_summary), cudaMemcpyDeviceToDevice), "restore");
  return B2BA_OK;
}

int64_t put(b2ba* h, const void* src, size_t bytes, void* dst, int64_t cap) {
  if ((int64_t)bytes > cap) return fail(h, B2BA_E_INVALID, "dump: destination too small");
  std::memcpy(dst, src, bytes);
  return (int64_t)bytes;
}

}  // namespace

// pt_ptr (point_major) or cam_ptr and the landmark schedule on the host, for the canonical index dumps
static int canonical_inputs(b2ba* h, bool point_major, std::vector<int>& ptr, std::vector<int>& sched) {
  ptr.resize((point_major ? h->np_tot : h->nc_tot) + 1);
  sched.resize(h->np_tot);
  CU(cudaMemcpy(ptr.data(), point_major ? h->d_pt_ptr.p : h->d_cam_ptr.p, ptr.size() * 4, cudaMemcpyDeviceToHost), "read");
  if (h->np_tot) CU(cudaMemcpy(sched.data(), h->d_sched.p, sched.size() * 4, cudaMemcpyDeviceToHost), "read");
  return 0;
}

int64_t b2ba_dump(b2ba_t* h, int what, void* dst, int64_t bytes) {
  const bool pose_hist = what >= B2BA_D_POSE_HIST_COST && what <= B2BA_D_POSE_HIST_ACCEPT;   // needs no problem
  if (!h || !dst || (h->path == PATH_NONE && !pose_hist)) return fail(h, B2BA_E_STATE, "no problem set");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  CU(cudaDeviceSynchronize(), "sync");
  const size_t no = (size_t)h->no;
  switch (what) {
    case B2BA_D_PM_PERM:
    case B2BA_D_CM_PERM: {
      std::vector<ObsRec> rec(no);
      CU(cudaMemcpy(rec.data(), what == B2BA_D_PM_PERM ? h->d_pm.p : h->d_cm.p, no * sizeof(ObsRec), cudaMemcpyDeviceToHost), "read");
      std::vector<int> perm(no);
      int b = 0;
      for (size_t a = 0; a < no; ++a) {
        while ((long long)a >= h->h_obs_off[b + 1]) ++b;
        perm[a] = rec[a].orig - (int)h->h_obs_off[b];
      }
      if (h->sched_order) {
        // the engine keeps the streams in landmark-schedule order; the dump is the CANONICAL order of
        // ba_index.cuh (point-major by landmark id, camera-major by (camera, landmark id)), rebuilt here
        std::vector<int> ptr, sched;
        if (int rc = canonical_inputs(h, what == B2BA_D_PM_PERM, ptr, sched)) return rc;
        std::vector<int> out(no);
        if (what == B2BA_D_PM_PERM) {
          size_t o = 0;
          for (int w = 0; w < h->B; ++w) {
            const int p0 = h->h_pt_off[w], npw = h->h_pt_off[w + 1] - p0;
            std::vector<int> pos(npw);
            for (int q = 0; q < npw; ++q) pos[sched[p0 + q]] = q;
            for (int p = 0; p < npw; ++p)
              for (int a = ptr[p0 + pos[p]]; a < ptr[p0 + pos[p] + 1]; ++a) out[o++] = perm[a];
          }
        } else {
          int w = 0;
          for (int c = 0; c < h->nc_tot; ++c) {
            while (c >= h->h_cam_off[w + 1]) ++w;
            std::vector<std::pair<int, int>> seg;   // (landmark id, original observation): rec.idx is a slot
            for (int a = ptr[c]; a < ptr[c + 1]; ++a) seg.emplace_back(sched[h->h_pt_off[w] + rec[a].idx], perm[a]);
            std::sort(seg.begin(), seg.end());
            for (int a = ptr[c]; a < ptr[c + 1]; ++a) out[a] = seg[a - ptr[c]].second;
          }
        }
        return put(h, out.data(), no * 4, dst, bytes);
      }
      return put(h, perm.data(), no * 4, dst, bytes);
    }
    case B2BA_D_PT_PTR:
    case B2BA_D_CAM_PTR: {
      const size_t n = (what == B2BA_D_PT_PTR ? h->np_tot : h->nc_tot) + 1;
      std::vector<int> v(n);
      CU(cudaMemcpy(v.data(), what == B2BA_D_PT_PTR ? h->d_pt_ptr.p : h->d_cam_ptr.p, n * 4, cudaMemcpyDeviceToHost), "read");
      if (what == B2BA_D_PT_PTR && h->sched_order) {
        std::vector<int> ptr, sched, out(n);
        if (int rc = canonical_inputs(h, true, ptr, sched)) return rc;
        int acc = 0;
        for (int w = 0; w < h->B; ++w) {
          const int p0 = h->h_pt_off[w], npw = h->h_pt_off[w + 1] - p0;
          std::vector<int> pos(npw);
          for (int q = 0; q < npw; ++q) pos[sched[p0 + q]] = q;
          for (int p = 0; p < npw; ++p) { out[p0 + p] = acc; acc += ptr[p0 + pos[p] + 1] - ptr[p0 + pos[p]]; }
        }
        out[h->np_tot] = acc;
        return put(h, out.data(), n * 4, dst, bytes);
      }
      return put(h, v.data(), n * 4, dst, bytes);
    }
    case B2BA_D_SCHED: {
      std::vector<int> v(h->np_tot);
      CU(cudaMemcpy(v.data(), h->d_sched.p, v.size() * 4, cudaMemcpyDeviceToHost), "read");
      return put(h, v.data(), v.size() * 4, dst, bytes);
    }
    case B2BA_D_FIDX: {
      std::vector<int> v(h->nc_tot);
      for (int b = 0; b < h->B; ++b) {
        int nf = 0;
        for (int c = h->h_cam_off[b]; c < h->h_cam_off[b + 1]; ++c) v[c] = h->h_fixed[c] ? -1 : nf++;
      }
      return put(h, v.data(), v.size() * 4, dst, bytes);
    }
    case B2BA_D_RESIDUALS: case B2BA_D_WEIGHTS: case B2BA_D_JC: case B2BA_D_JP: {
      const size_t per = what == B2BA_D_RESIDUALS ? 2 : what == B2BA_D_WEIGHTS ? 1 : what == B2BA_D_JC ? 12 : 6;
      DevBuf<double> tmp;
      CU(tmp.alloc(no * per), "alloc");
      const double* pts_lm = h->d_pts0.p;
      if (h->sched_order) {
        const int rcp = permute_pts(h, h->d_pts0.p, h->d_pts1.p, false, nullptr);
        if (rcp) return cuda_fail(h, (cudaError_t)rcp, "permute points");
        pts_lm = h->d_pts1.p;
      }
      int rc = dump_obs_terms(make_index_in(h), make_params(h, 1), h->d_poses.p, pts_lm,
                              what == B2BA_D_RESIDUALS ? tmp.p : nullptr, what == B2BA_D_WEIGHTS ? tmp.p : nullptr,
                              what == B2BA_D_JC ? tmp.p : nullptr, what == B2BA_D_JP ? tmp.p : nullptr, nullptr);
      if (rc) return cuda_fail(h, (cudaError_t)rc, "dump_obs_terms");
      if ((int64_t)(no * per * 8) > bytes) return fail(h, B2BA_E_INVALID, "dump: destination too small");
      CU(cudaMemcpy(dst, tmp.p, no * per * 8, cudaMemcpyDeviceToHost), "read");
      return (int64_t)(no * per * 8);
    }
    case B2BA_D_BLK_PTR:
    case B2BA_D_BLK_COL: {
      if (h->path != PATH_LARGE) return fail(h, B2BA_E_UNSUPPORTED, "S block structure exists on the large (PCG) path only");
      const size_t n = what == B2BA_D_BLK_PTR ? (size_t)h->lix.nfree + 1 : (size_t)h->lix.nblk;
      std::vector<int> v(n);
      CU(cudaMemcpy(v.data(), what == B2BA_D_BLK_PTR ? h->lix.blk_ptr.p : h->lix.blk_col.p, n * 4, cudaMemcpyDeviceToHost), "read");
      return put(h, v.data(), n * 4, dst, bytes);
    }
    case B2BA_D_B: case B2BA_D_GC: case B2BA_D_C: case B2BA_D_GP: case B2BA_D_S: case B2BA_D_GRED:
    case B2BA_D_DC: case B2BA_D_DP: case B2BA_D_SCALARS: {
      if (h->path == PATH_LARGE) {
        if (what != B2BA_D_S && what != B2BA_D_GRED && what != B2BA_D_DC && what != B2BA_D_SCALARS)
          return fail(h, B2BA_E_UNSUPPORTED, "large path dumps S, g_red, delta_c and the scalars only");
        if (h->nranks > 1) return fail(h, B2BA_E_UNSUPPORTED, "stage dumps are single-GPU");
        DevBuf<double> sp, sx;
        DevBuf<b2ba_summary> ss;
        const size_t nc12 = (size_t)h->nc_tot * 12, np3 = (size_t)h->np_tot * 3, n = 6 * (size_t)h->lix.nfree;
        CU(sp.alloc(nc12), "alloc"); CU(sx.alloc(np3), "alloc"); CU(ss.alloc(1), "alloc");
        CU(cudaMemcpy(sp.p, h->d_poses.p, nc12 * 8, cudaMemcpyDeviceToDevice), "save");
        CU(cudaMemcpy(sx.p, h->d_pts0.p, np3 * 8, cudaMemcpyDeviceToDevice), "save");
        CU(cudaMemcpy(ss.p, h->d_summ.p, sizeof(b2ba_summary), cudaMemcpyDeviceToDevice), "save");
        int rc = solve_large(h, make_params(h, 1), false, nullptr);
        if (rc) return rc;
        CU(cudaDeviceSynchronize(), "stage");
        std::vector<double> sg((size_t)h->lix.nblk * 36 + n), dcv(n);
        std::vector<int> bp(h->lix.nfree + 1), bc(h->lix.nblk);
        LargeCtrl ctrl;
        CU(cudaMemcpy(sg.data(), h->l_Sg.p, sg.size() * 8, cudaMemcpyDeviceToHost), "read");
        CU(cudaMemcpy(dcv.data(), h->l_dc.p, n * 8, cudaMemcpyDeviceToHost), "read");
        CU(cudaMemcpy(bp.data(), h->lix.blk_ptr.p, bp.size() * 4, cudaMemcpyDeviceToHost), "read");
        CU(cudaMemcpy(bc.data(), h->lix.blk_col.p, bc.size() * 4, cudaMemcpyDeviceToHost), "read");
        CU(cudaMemcpy(&ctrl, h->l_ctrl.p, sizeof(ctrl), cudaMemcpyDeviceToHost), "read");
        CU(cudaMemcpy(h->d_poses.p, sp.p, nc12 * 8, cudaMemcpyDeviceToDevice), "restore");
        CU(cudaMemcpy(h->d_pts0.p, sx.p, np3 * 8, cudaMemcpyDeviceToDevice), "restore");
        CU(cudaMemcpy(h->d_summ.p, ss.p, sizeof(b2ba_summary), cudaMemcpyDeviceToDevice), "restore");
        if (what == B2BA_D_GRED) return put(h, &sg[(size_t)h->lix.nblk * 36], n * 8, dst, bytes);
        if (what == B2BA_D_DC) return put(h, dcv.data(), n * 8, dst, bytes);
        if (what == B2BA_D_SCALARS) {
          const double sc[4] = {ctrl.initial_cost, 0.0, ctrl.accept ? ctrl.cost : 0.0, (double)ctrl.pcg_last};
          return put(h, sc, sizeof(sc), dst, bytes);
        }
        std::vector<double> S(n * n, 0.0);
        for (int fi = 0; fi < h->lix.nfree; ++fi)
          for (int b = bp[fi]; b < bp[fi + 1]; ++b) {
            const int fj = bc[b];
            for (int r = 0; r < 6; ++r)
              for (int c = 0; c < 6; ++c) {
                if (fi == fj && c < r) continue;
                const double v = sg[(size_t)b * 36 + 6 * r + c];
                S[(size_t)(6 * fi + r) * n + 6 * fj + c] = v;
                S[(size_t)(6 * fj + c) * n + 6 * fi + r] = v;
              }
          }
        return put(h, S.data(), n * n * 8, dst, bytes);
      }
      std::vector<double> o;
      WinDbg L;
      int rc = run_stage(h, o, L);
      if (rc) return rc;
      const size_t n = L.n, nc = L.nc, np = L.np;
      switch (what) {
        case B2BA_D_B: return put(h, &o[L.off_B()], nc * 36 * 8, dst, bytes);
        case B2BA_D_GC: return put(h, &o[L.off_gc()], nc * 6 * 8, dst, bytes);
        case B2BA_D_C: case B2BA_D_GP: case B2BA_D_DP: {
          const size_t per = what == B2BA_D_C ? 6 : 3;
          const double* src = &o[what == B2BA_D_C ? L.off_C() : what == B2BA_D_GP ? L.off_gp() : L.off_dp()];
          if (!h->sched_order) return put(h, src, np * per * 8, dst, bytes);
          std::vector<int> ptr, sched;       // single window: slot q holds landmark sched[q]
          if (int rcs = canonical_inputs(h, true, ptr, sched)) return rcs;
          std::vector<double> lm(np * per);
          for (size_t q = 0; q < np; ++q)
            for (size_t e = 0; e < per; ++e) lm[(size_t)sched[q] * per + e] = src[q * per + e];
          return put(h, lm.data(), np * per * 8, dst, bytes);
        }
        case B2BA_D_S: return put(h, &o[L.off_S()], n * n * 8, dst, bytes);
        case B2BA_D_GRED: return put(h, &o[L.off_gred()], n * 8, dst, bytes);
        case B2BA_D_DC: return put(h, &o[L.off_dc()], n * 8, dst, bytes);
        default: return put(h, &o[L.off_scal()], 4 * 8, dst, bytes);
      }
    }
    case B2BA_D_POSE_HIST_COST: case B2BA_D_POSE_HIST_MU: case B2BA_D_POSE_HIST_ACCEPT: {
      if (!h->r_nb || !(h->opt.flags & 1)) return fail(h, B2BA_E_STATE, "pose history not recorded (options.flags bit 0)");
      const size_t n = (size_t)h->r_nb * kPoseHist, sz = what == B2BA_D_POSE_HIST_ACCEPT ? 4 : 8;
      if ((int64_t)(n * sz) > bytes) return fail(h, B2BA_E_INVALID, "dump: destination too small");
      const void* src = what == B2BA_D_POSE_HIST_COST ? (void*)h->r_hist_cost.p
                        : what == B2BA_D_POSE_HIST_MU ? (void*)h->r_hist_mu.p : (void*)h->r_hist_accept.p;
      CU(cudaMemcpy(dst, src, n * sz, cudaMemcpyDeviceToHost), "read");
      return (int64_t)(n * sz);
    }
    case B2BA_D_HIST_COST: case B2BA_D_HIST_MU: case B2BA_D_HIST_ACCEPT: {
      if (!h->solved || !(h->opt.flags & 1)) return fail(h, B2BA_E_STATE, "history not recorded (options.flags bit 0)");
      const size_t n = (size_t)h->B * kHist, sz = what == B2BA_D_HIST_ACCEPT ? 4 : 8;
      if ((int64_t)(n * sz) > bytes) return fail(h, B2BA_E_INVALID, "dump: destination too small");
      const void* src = what == B2BA_D_HIST_COST ? (void*)h->d_hist_cost.p
                        : what == B2BA_D_HIST_MU ? (void*)h->d_hist_mu.p : (void*)h->d_hist_accept.p;
      CU(cudaMemcpy(dst, src, n * sz, cudaMemcpyDeviceToHost), "read");
      return (int64_t)(n * sz);
    }
    default:
      return fail(h, B2BA_E_UNSUPPORTED, "dump: unknown or unavailable array");
  }
}

// ---------------------------------------------------------------------------------------------
// b2ba_solve_window_noalloc: the window kernel with its scratch in a file-scope device array (include/b2ba.h).
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// Scratch layout, per landmark of the handle's problem (schedule order), 18 doubles:
//   [0, 3 Np)  landmarks, current      [3 Np, 6 Np)  landmarks, trial (ping-pong)      [6 Np, 18 Np)  F | g_p | D_p rows
// Every device that loads the library holds the array (kNoAllocMaxPoints * 18 * 8 bytes = 4.5 MiB) until exit.
// In HOST code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address: the address a kernel
// may use comes from cudaGetSymbolAddress, looked up once per device with that device current.
constexpr int kNoAllocMaxPoints = B2BA_NOALLOC_MAX_POINTS;
constexpr size_t kNoAllocDoublesPerPoint = 6 + kPtScratch;
static __device__ __align__(256) double g_noalloc_scratch[(size_t)kNoAllocMaxPoints * kNoAllocDoublesPerPoint];

namespace {
// device address of g_noalloc_scratch on the CURRENT device (resolved once per device, under a mutex)
int noalloc_scratch_ptr(b2ba* h, double** out) {
  static std::mutex mtx;
  static double* cache[64] = {};
  int dev = -1;
  CU(cudaGetDevice(&dev), "cudaGetDevice");
  if (dev < 0 || dev >= 64) return fail(h, B2BA_E_UNSUPPORTED, "solve_window_noalloc: device index above 63");
  std::lock_guard<std::mutex> lock(mtx);
  if (!cache[dev]) {
    double* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) return cuda_fail(h, e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    cache[dev] = ptr;
  }
  *out = cache[dev];
  return B2BA_OK;
}
}  // namespace

int b2ba_solve_window_noalloc(b2ba_t* h, double* poses_cw, double* points, b2ba_summary* summaries, int32_t max_iters,
                              void* cuda_stream) {
  if (!h) return B2BA_E_INVALID;
  if (h->path != PATH_WIN_SMALL && h->path != PATH_WIN_BIG)
    return fail(h, B2BA_E_STATE, "solve_window_noalloc: set a window problem first (b2ba_set_problem / _batch; <= 24 cameras per window)");
  if (!poses_cw || !points || !summaries) return fail(h, B2BA_E_INVALID, "solve_window_noalloc: null argument (device pointers required)");
  if (h->np_tot > kNoAllocMaxPoints)
    return fail(h, B2BA_E_UNSUPPORTED, "solve_window_noalloc: " + std::to_string(h->np_tot) + " landmarks exceed the static scratch (B2BA_NOALLOC_MAX_POINTS = " +
                                       std::to_string(kNoAllocMaxPoints) + "); use b2ba_solve");
  DeviceGuard dg_(h->device); CU(dg_.err, "cudaSetDevice");
  double* scr = nullptr;
  if (const int rc = noalloc_scratch_ptr(h, &scr)) return rc;
  cudaStream_t s = (cudaStream_t)cuda_stream;
  const size_t np = (size_t)h->np_tot;
  // the handle's index arrays (built by set_problem) + the caller's buffers + the static scratch; nothing of the
  // handle's own state (d_poses, d_pts*, d_scratch, summaries, history) is read or written
  WinData d = make_win(h, false, nullptr);
  d.poses = poses_cw;
  d.pts[0] = scr; d.pts[1] = scr + 3 * np; d.pt_scratch = scr + 6 * np;
  d.summaries = summaries;
  // the array keeps the last call's contents.  Every word the kernel reads is written first in the same call (landmarks by
  // the permutation below, trial landmarks and scratch rows by the kernel before it reads them); the scratch rows are
  // cleared all the same, so that a call never depends on what an earlier one left behind
  CU(cudaMemsetAsync(scr + 3 * np, 0, (size_t)(3 + kPtScratch) * np * sizeof(double), s), "clear scratch");
  int rc = h->sched_order ? permute_pts(h, points, d.pts[0], true, s)
                          : (int)cudaMemcpyAsync(d.pts[0], points, 3 * np * sizeof(double), cudaMemcpyDeviceToDevice, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "solve_window_noalloc: landmarks in");
  rc = launch_window_path(h, d, make_params(h, max_iters), s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "solve_window_noalloc: launch");
  rc = h->sched_order ? permute_pts(h, d.pts[0], points, false, s)
                      : (int)cudaMemcpyAsync(points, d.pts[0], 3 * np * sizeof(double), cudaMemcpyDeviceToDevice, s);
  if (rc) return cuda_fail(h, (cudaError_t)rc, "solve_window_noalloc: landmarks out");
  return B2BA_OK;      // enqueued on `cuda_stream`; no synchronisation, no allocation
}

}  // extern "C"
