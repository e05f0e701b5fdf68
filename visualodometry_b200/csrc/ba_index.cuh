// a8: observation-to-block index builder, on the device, once per problem.
// Canonical orders (the bit-exact contract the parity tests check):
//   point-major : observations sorted by (window, pt, cam, original index)
//   camera-major: observations sorted by (window, cam, pt, original index)
#pragma once
#include "ba_math.cuh"

namespace b2 {

struct IndexIn {
  int nb, nc_tot, np_tot;
  long long no;
  const int* cam_off;        // device, nb+1
  const int* pt_off;         // device, nb+1
  const long long* obs_off;  // device, nb+1
  const int* obs_cam;        // device, window-local
  const int* obs_pt;         // device, window-local
  const float2* obs_uv;      // device
  int max_np = 0, max_nc = 0;   // host-known maxima over the windows (0 = unknown: the global-sort builder is used)
};

// window of observation i: obs_off[b] <= i < obs_off[b + 1]
__device__ __forceinline__ int find_window(const long long* __restrict__ obs_off, int nb, long long i) {
  int lo = 0, hi = nb;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (obs_off[mid] <= i) lo = mid; else hi = mid;
  }
  return lo;
}

struct IndexOut {
  ObsRec* pm;      // no
  int* pt_ptr;     // np_tot+1
  ObsRec* cm;      // no
  int* cam_ptr;    // nc_tot+1
  int* flags;      // [0] out-of-range index, [1] duplicate (cam,pt) observation
  int* sched;      // np_tot or nullptr: per window, the window-local landmark ids ordered by decreasing
                   // observation count (stable).  When given, the whole point side is laid out in this order
                   // (slot q of a window = landmark sched[q]): pm, pt_ptr, the landmark index of cm records;
                   // the landmarks sharing a warp of the window kernel then have (nearly) the same number of
                   // observations and every per-landmark array is walked contiguously.
};

// Builds both streams. Returns cudaError_t as int. Temporary storage is allocated and freed
// internally (stream-ordered); the call is asynchronous on `s`.
int build_index(const IndexIn& in, const IndexOut& out, cudaStream_t s);

// landmark rows (3 doubles) between landmark order and schedule order (src != dst)
int permute_points(int nb, int np_tot, const int* pt_off, const int* sched, const double* src, double* dst,
                   bool to_sched, cudaStream_t s);

// reprojection error (pixels) per observation in ORIGINAL observation order from landmark-ordered points;
// -1 where the landmark is behind the cheirality gate (the input of outlier culling)
int obs_errors(const IndexIn& in, const Params& P, const double* poses, const double* pts, double* err,
               cudaStream_t s);

// SURVEY §8 row f-4, outlier culling after BA, in ONE kernel (ba_cull.cu): reprojection error of every observation at
// the given state -> keep = (0 <= err <= max_err), the reference's landmark gate (frontend.py:139-142, config.py:31-32)
// -> stable compaction of the observation arrays (single-pass scan with decoupled look-back) + surviving observations
// per landmark (`live`, np_tot) and per window (`win_kept`, nb) + the total (`n_kept`).  Any output may be null.
// `work`: device scratch of cull_work_bytes(no) bytes.  Asynchronous on `s`.
size_t cull_work_bytes(long long no);
int cull_observations(const IndexIn& in, const Params& P, const double* poses, const double* pts, double max_err,
                      unsigned char* keep, int* live, long long* win_kept, int* out_cam, int* out_pt, float2* out_uv,
                      int* kept_index, long long* n_kept, void* work, cudaStream_t s);

// a1 dump in ORIGINAL observation order (parity tests): r (no x 2), w (no), Jc (no x 12), Jp (no x 6)
int dump_obs_terms(const IndexIn& in, const Params& P, const double* poses, const double* pts,
                   double* r, double* w, double* jc, double* jp, cudaStream_t s);

}  // namespace b2
