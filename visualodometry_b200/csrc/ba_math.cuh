// Device math shared by every b2ba kernel: pinhole residual + robust weight + analytic
// Jacobians, 3x3 SPD factor (a3), SO(3) retraction (a7).
// Conventions follow the reference: x_c = R_cw X + t_cw, uv = K x_c / z, zero distortion
// (src/modules/frontend.py:128-139, src/modules/vo.py:139), residual = projected - observed in
// pixels (frontend.py:140), cheirality gate z >= z_min (frontend.py:135, config.py:31).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

constexpr double kDiagMin = 1e-6, kDiagMax = 1e32, kMinRatio = 1e-3, kMuMax = 1e16, kMuMin = 1e-32;

struct Params {     // per-solve constants, passed by value to kernels
  double fx, fy, cx, cy;
  double z_min, huber_a, init_mu, ftol, gtol, ptol;
  int loss, max_iters;
};

// One observation record of the SoA streams: 16 B, read with a single LDG.128.
//   point-major stream : idx = window-local camera, orig = original observation index
//   camera-major stream: idx = window-local point
struct __align__(16) ObsRec { float u, v; int idx; int orig; };

__device__ __forceinline__ double clampd(double v) { return fmin(fmax(v, kDiagMin), kDiagMax); }

struct Proj {      // camera-frame point and residual
  double x, y, z, iz, r0, r1, w, rho;
  bool valid;
};

// residual + robust weight. pose: 12 doubles [R|t] (shared or global memory).
__device__ __forceinline__ void project(const Params& P, const double* __restrict__ pose, double X0,
                                        double X1, double X2, float u, float v, Proj& o) {
  o.x = fma(pose[0], X0, fma(pose[1], X1, fma(pose[2], X2, pose[9])));
  o.y = fma(pose[3], X0, fma(pose[4], X1, fma(pose[5], X2, pose[10])));
  o.z = fma(pose[6], X0, fma(pose[7], X1, fma(pose[8], X2, pose[11])));
  o.valid = o.z >= P.z_min;
  o.iz = 1.0 / o.z;
  o.r0 = fma(P.fx * o.x, o.iz, P.cx - (double)u);
  o.r1 = fma(P.fy * o.y, o.iz, P.cy - (double)v);
  double s = o.r0 * o.r0 + o.r1 * o.r1;
  o.w = 1.0;
  o.rho = s;
  if (P.loss == 1 && s > P.huber_a * P.huber_a) {
    double rs = sqrt(s);
    o.rho = 2.0 * P.huber_a * rs - P.huber_a * P.huber_a;
    o.w = P.huber_a / rs;
  }
  if (!o.valid) { o.r0 = 0; o.r1 = 0; o.w = 0; o.rho = 0; o.iz = 0; }
}

// J_p = dr/dx * R  (2x3, row-major in jp[6]);  dr/dx = [[a0,0,a2],[0,b1,b2]]
__device__ __forceinline__ void jac_point(const Params& P, const double* __restrict__ pose,
                                          const Proj& o, double* jp) {
  double a0 = P.fx * o.iz, a2 = -P.fx * o.x * o.iz * o.iz;
  double b1 = P.fy * o.iz, b2 = -P.fy * o.y * o.iz * o.iz;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    jp[j] = a0 * pose[j] + a2 * pose[6 + j];
    jp[3 + j] = b1 * pose[3 + j] + b2 * pose[6 + j];
  }
}

// J_c = dr/dx * [ -[x]x | I ]  (2x6, row-major in jc[12]); left perturbation, ordering [dw, dt]
__device__ __forceinline__ void jac_cam(const Params& P, const Proj& o, double* jc) {
  double a0 = P.fx * o.iz, a2 = -P.fx * o.x * o.iz * o.iz;
  double b1 = P.fy * o.iz, b2 = -P.fy * o.y * o.iz * o.iz;
  jc[0] = a2 * o.y;             jc[1] = a0 * o.z - a2 * o.x;  jc[2] = -a0 * o.y;
  jc[3] = a0;                   jc[4] = 0.0;                  jc[5] = a2;
  jc[6] = -b1 * o.z + b2 * o.y; jc[7] = -b2 * o.x;            jc[8] = b1 * o.x;
  jc[9] = 0.0;                  jc[10] = b1;                  jc[11] = b2;
}

// E = w J_c^T J_p (6x3 row-major)
__device__ __forceinline__ void make_E(double w, const double* jc, const double* jp, double* E) {
#pragma unroll
  for (int r = 0; r < 6; ++r)
#pragma unroll
    for (int j = 0; j < 3; ++j) E[3 * r + j] = w * (jc[r] * jp[j] + jc[6 + r] * jp[3 + j]);
}

// Upper-triangular F (f00 f01 f02 f11 f12 f22) with C'^-1 = F F^T, C' SPD 3x3 given as upper
// (c00 c01 c02 c11 c12 c22).  F = G^-T for the Cholesky factor C' = G G^T.  false if not SPD.
__device__ __forceinline__ bool sym3_inv_factor(const double* C, double* F) {
  double a = C[0];
  if (!(a > 0.0)) return false;
  a = sqrt(a);
  double ia = 1.0 / a;
  double b = C[1] * ia, d = C[2] * ia;
  double c2 = C[3] - b * b;
  if (!(c2 > 0.0)) return false;
  double c = sqrt(c2), ic = 1.0 / c;
  double e = (C[4] - b * d) * ic;
  double f2 = C[5] - d * d - e * e;
  if (!(f2 > 0.0) || !isfinite(f2)) return false;
  double f = sqrt(f2), ifr = 1.0 / f;
  // G^-1 = [[ia,0,0],[-b ia ic, ic, 0],[(b e - c d) ia ic if, -e ic if, if]] ; F = (G^-1)^T
  F[0] = ia;  F[1] = -b * ia * ic;  F[2] = (b * e - c * d) * ia * ic * ifr;
  F[3] = ic;  F[4] = -e * ic * ifr;
  F[5] = ifr;
  return true;
}

// out = Exp(d[0..2]) * pose, t = Exp(dw) t + d[3..5]   (Spec a0 retraction)
__device__ __forceinline__ void retract_pose(const double* pose, const double* d, double* out) {
  double x = d[0], y = d[1], z = d[2];
  double t2 = x * x + y * y + z * z, a, b;
  if (t2 < 1e-16) { a = 1.0 - t2 / 6.0; b = 0.5 - t2 / 24.0; }
  else { double t = sqrt(t2); a = sin(t) / t; b = (1.0 - cos(t)) / t2; }
  double E[9];
  E[0] = 1 + b * (-(y * y + z * z)); E[1] = -a * z + b * x * y;         E[2] = a * y + b * x * z;
  E[3] = a * z + b * x * y;          E[4] = 1 + b * (-(x * x + z * z)); E[5] = -a * x + b * y * z;
  E[6] = -a * y + b * x * z;         E[7] = a * x + b * y * z;          E[8] = 1 + b * (-(x * x + y * y));
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j)
      out[3 * i + j] = E[3 * i] * pose[j] + E[3 * i + 1] * pose[3 + j] + E[3 * i + 2] * pose[6 + j];
    out[9 + i] = E[3 * i] * pose[9] + E[3 * i + 1] * pose[10] + E[3 * i + 2] * pose[11] + d[3 + i];
  }
}

// ---------------------------------------------------------------------------------------------
// Lean forms used by the window kernel's hot loops.  Same quantities as project / jac_cam /
// jac_point / make_E above, written as explicit FMA chains that skip the structural zeros of
// J_c (J_c[0][4] = J_c[1][3] = 0); differences to the forms above are pure rounding (~1e-16).
struct Pose12 { double r[12]; };   // [R row-major | t] in registers

__device__ __forceinline__ void load_pose(const double* __restrict__ p, Pose12& q) {   // p: 16-byte aligned
  const double2* v = reinterpret_cast<const double2*>(p);
#pragma unroll
  for (int i = 0; i < 6; ++i) { const double2 t = v[i]; q.r[2 * i] = t.x; q.r[2 * i + 1] = t.y; }
}

struct Geo {       // one observation: camera-frame point, residual, weight, dr/dx = [[a0,0,a2],[0,b1,b2]]
  double x, y, z, r0, r1, w, rho, a0, a2, b1, b2;
  bool valid;
};

// 1/z without the slow-path branch of the IEEE division: rcp.approx.ftz.f64 seed (MUFU.RCP64H, ~2^-20)
// + two Newton steps in fp64 (-> below 2^-52 relative; not correctly rounded, 1-2 ulp).  z >= z_min here.
__device__ __forceinline__ double rcp_depth(double z) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(z));
  y = y * fma(-z, y, 2.0);
  y = y * fma(-z, y, 2.0);
  return y;
}

// 1/sqrt(x) for x > 0 without the library's slow path: rsqrt.approx.ftz.f64 seed (MUFU.RSQ64H, ~2^-20) + ONE third-order
// step y (1 + e/2 + 3 e^2 / 8), e = 1 - x y^2 (error ~ e^3: below 2^-52 relative).  Branch-free; the dependent chain is
// four FP64 operations after the seed (two Newton steps were six).
__device__ __forceinline__ double rsqrt_pos(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-(x * y), y, 1.0);
  const double p = fma(0.375, e, 0.5);
  return fma(y * e, p, y);
}

// sym3_inv_factor with reciprocal square roots only (no sqrt, no division): F = G^-T for C' = G G^T.
__device__ __forceinline__ bool sym3_inv_factor_fast(const double* C, double* F) {
  const double a = C[0];
  const double ia = rsqrt_pos(a);
  const double b = C[1] * ia, d = C[2] * ia;
  const double c2 = fma(-b, b, C[3]);
  const double ic = rsqrt_pos(c2);
  const double e = fma(-b, d, C[4]) * ic;
  const double f2 = fma(-e, e, fma(-d, d, C[5]));
  const double ifr = rsqrt_pos(f2);
  const double c = c2 * ic;                       // sqrt(c2)
  F[0] = ia;  F[1] = -b * ia * ic;  F[2] = fma(b, e, -(c * d)) * ia * ic * ifr;
  F[3] = ic;  F[4] = -e * ic * ifr;
  F[5] = ifr;
  return (a > 0.0) && (c2 > 0.0) && (f2 > 0.0) && isfinite(f2);
}

// fire-and-forget L1 prefetch of the 128-byte line holding p (no register cost)
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// LOSS is compile-time in the window kernel (0 trivial, 1 Huber): the trivial path is branch-free, so
// two observations handled by one lane interleave in the instruction stream.
template <int LOSS>
__device__ __forceinline__ void geo_project(const Params& P, const Pose12& q, double X0, double X1, double X2,
                                            float u, float v, Geo& o) {
  o.x = fma(q.r[0], X0, fma(q.r[1], X1, fma(q.r[2], X2, q.r[9])));
  o.y = fma(q.r[3], X0, fma(q.r[4], X1, fma(q.r[5], X2, q.r[10])));
  o.z = fma(q.r[6], X0, fma(q.r[7], X1, fma(q.r[8], X2, q.r[11])));
  o.valid = o.z >= P.z_min;
  const double iz = o.valid ? rcp_depth(o.z) : 0.0;
  o.a0 = P.fx * iz;
  o.b1 = P.fy * iz;
  o.r0 = fma(o.a0, o.x, P.cx - (double)u);
  o.r1 = fma(o.b1, o.y, P.cy - (double)v);
  o.a2 = -(o.a0 * o.x) * iz;
  o.b2 = -(o.b1 * o.y) * iz;
  const double s = fma(o.r0, o.r0, o.r1 * o.r1);
  o.w = 1.0;
  o.rho = s;
  if (LOSS == 1) {
    if (s > P.huber_a * P.huber_a) {
      const double rs = sqrt(s);
      o.rho = 2.0 * P.huber_a * rs - P.huber_a * P.huber_a;
      o.w = P.huber_a / rs;
    }
  }
  if (!o.valid) { o.r0 = 0; o.r1 = 0; o.rho = 0; o.w = 0; }
}

// J_p rows: jp[0..2] = a0 R0 + a2 R2, jp[3..5] = b1 R1 + b2 R2
__device__ __forceinline__ void geo_jp(const Pose12& q, const Geo& o, double* jp) {
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    jp[j] = fma(o.a0, q.r[j], o.a2 * q.r[6 + j]);
    jp[3 + j] = fma(o.b1, q.r[3 + j], o.b2 * q.r[6 + j]);
  }
}

// J_c rows (left perturbation [dw, dt]); c0[4] and c1[3] are structurally zero and never read
__device__ __forceinline__ void geo_jc(const Geo& o, double* c0, double* c1) {
  c0[0] = o.a2 * o.y;                 c0[1] = fma(o.a0, o.z, -(o.a2 * o.x)); c0[2] = -(o.a0 * o.y);
  c0[3] = o.a0;                       c0[4] = 0.0;                            c0[5] = o.a2;
  c1[0] = fma(o.b2, o.y, -(o.b1 * o.z)); c1[1] = -(o.b2 * o.x);               c1[2] = o.b1 * o.x;
  c1[3] = 0.0;                        c1[4] = o.b1;                           c1[5] = o.b2;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// D (8x8) += A (8x4) * B (4x8), fp64 tensor-core MMA (DMMA).  Fragment layout (PTX ISA,
// mma.m8n8k4.f64): a = A[lane/4][lane%4], b = B[lane%4][lane/4], c0/c1 = C[lane/4][2*(lane%4)+{0,1}]
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

}  // namespace b2
