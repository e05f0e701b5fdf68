// Motion-only BA (pose refinement against fixed landmarks), SURVEY §8 row f-2 -- see ba_pose.cu.
#pragma once
#include "ba_math.cuh"
#include "../../include/b2ba.h"

namespace b2 {

constexpr int kPoseHist = 64;

struct PoseData {
  int nb;                      // pose problems
  const long long* obs_off;    // nb+1 (device): problem b owns observations [obs_off[b], obs_off[b+1])
  double* poses;               // nb x 12 [R row-major | t] world->camera, in/out
  const double* points;        // No x 3: the landmark of every observation (fixed)
  const float2* obs_uv;        // No
  b2ba_summary* summaries;     // nb
  double* hist_cost;           // nb x kPoseHist or null
  double* hist_mu;
  int* hist_accept;
};

int launch_pose_lm(const PoseData& d, const Params& P, cudaStream_t s);
// robust tracking of ONE pose (d.nb == 1): rounds of masked LM + inlier re-classification, see ba_pose.cu
// classify_first: the mask starts as the consensus set of the input pose at thr and every round runs at thr
int launch_pose_track(const PoseData& d, const Params& P, double thr, int rounds, unsigned char* mask,
                      int* n_inliers, cudaStream_t s, int classify_first = 0);

// sampled consensus: nh hypotheses, each the LM fit (hyp_iters iterations from the prior) of kRansacSample sampled
// observations; the one with the largest consensus at thr replaces track.poses (then launch_pose_track refines it)
constexpr int kRansacSample = 4;
struct RansacData {
  int nh, hyp_iters;
  long long n;
  unsigned long long seed;
  double* hp_pts;            // nh x kRansacSample x 3
  float2* hp_uv;             // nh x kRansacSample
  double* hp_poses;          // nh x 12
  long long* hp_off;         // nh + 1
  int* hp_idx;               // nh x kRansacSample: the sampled observation indices
  b2ba_summary* hp_summ;     // nh
  int* score;                // nh
  int* best;                 // [index, consensus] of the winning hypothesis
};
int launch_pose_ransac(const RansacData& r, const PoseData& track, const Params& P, double thr, cudaStream_t s);

// row f-3: two-view DLT triangulation + the reference's depth / reprojection gates (device arrays; P1, P2 on the host)
int launch_triangulate(int n, const double* P1, const double* P2, const float2* uv1, const float2* uv2,
                       double min_depth, double max_err, double* X, unsigned char* keep, cudaStream_t s);

}  // namespace b2
