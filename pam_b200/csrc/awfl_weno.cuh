// Device-side WENO reconstruction for the AWFL dycore (sm_100a).
//
// Mathematically identical to the reference's Dycore::reconstruct (dynamics/awfl/Dycore.h:591-604) +
// weno::compute_weno_coefs (dynamics/awfl/WenoLimiter.h:98-181), re-associated for the FP64 pipe:
//   * the "bridge" polynomial (WenoLimiter.h:131-138) is one mat-vec with a precomputed matrix
//     B = (sten_to_coefs - sum_i idl_i*embed(weno_lower_i)) / idl_m   (tools/gen_tables.py; per level on the host
//     for the vertical matrices), which removes ord divides and hs*hs FMAs per reconstruction;
//   * every division by a strictly positive quantity is a MUFU.RCP64H seed + two Newton steps (no slow path);
//   * the second convexify (WenoLimiter.h:166) is applied once to the blended edge values instead of to each weight;
//   * the blended polynomial is evaluated at BOTH cell edges from its even/odd parts, so one set of WENO weights per
//     cell serves the right state of its left interface and the left state of its right interface.
// All of this changes results at the 1e-16 relative level only (FP64 re-association), far inside the 1e-12 budget.
#pragma once
#include "awfl_tables.h"

namespace awfl {

template <int ORD> struct Cst;

#define AWFL_DEFINE_CST(O)                                                                                      \
  template <> struct Cst<O> {                                                                                   \
    static constexpr int HS = (O + 1) / 2;                                                                      \
    static __device__ __forceinline__ double lo(int i, int s, int ii) {                                         \
      constexpr double T[HS][HS][HS] = AWFL_WRL_##O##_INIT;                                                     \
      return T[i][s][ii];                                                                                       \
    }                                                                                                           \
    static __device__ __forceinline__ double br(int s, int ii) {                                                \
      constexpr double T[O][O] = AWFL_BRIDGE_##O##_INIT;                                                        \
      return T[s][ii];                                                                                          \
    }                                                                                                           \
    static __device__ __forceinline__ double idl(int i) {                                                       \
      constexpr double T[HS + 1] = AWFL_IDL_##O##_INIT;                                                         \
      return T[i];                                                                                              \
    }                                                                                                           \
    static __device__ __forceinline__ double sigma() { return AWFL_SIGMA_##O; }                                 \
    /* TransformMatrices.h coefs_to_tv for hs and ord coefficients (truncated squared-derivative integrals) */  \
    static __device__ __forceinline__ double tv_lo(const double (&a)[HS]) {                                     \
      constexpr int N = AWFL_TVLO_N_##O;                                                                        \
      constexpr int I[N] = AWFL_TVLO_I_##O##_INIT;                                                              \
      constexpr int J[N] = AWFL_TVLO_J_##O##_INIT;                                                              \
      constexpr double C[N] = AWFL_TVLO_C_##O##_INIT;                                                           \
      double r = (C[0] * a[I[0]]) * a[J[0]];                                                                    \
      _Pragma("unroll") for (int t = 1; t < N; t++) r = fma(C[t] * a[I[t]], a[J[t]], r);                        \
      return r;                                                                                                 \
    }                                                                                                           \
    static __device__ __forceinline__ double tv_hi(const double (&a)[O]) {                                      \
      constexpr int N = AWFL_TVHI_N_##O;                                                                        \
      constexpr int I[N] = AWFL_TVHI_I_##O##_INIT;                                                              \
      constexpr int J[N] = AWFL_TVHI_J_##O##_INIT;                                                              \
      constexpr double C[N] = AWFL_TVHI_C_##O##_INIT;                                                           \
      double r = (C[0] * a[I[0]]) * a[J[0]];                                                                    \
      _Pragma("unroll") for (int t = 1; t < N; t++) r = fma(C[t] * a[I[t]], a[J[t]], r);                        \
      return r;                                                                                                 \
    }                                                                                                           \
  };

AWFL_DEFINE_CST(3)
AWFL_DEFINE_CST(5)
AWFL_DEFINE_CST(7)
AWFL_DEFINE_CST(9)

// 1/x for x > 0, normal range: MUFU.RCP64H seed (relative error e0 ~ 2^-20) + one cubically convergent step
// r*(1 + e + e^2), e = 1 - x*r  (error e0^3 ~ 2^-60, then rounding): <= ~1 ulp in 3 DFMA, no slow path.
__device__ __forceinline__ double rcp_pos(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  const double e = fma(-x, r, 1.0);
  const double t = fma(e, e, e);
  return fma(r, t, r);
}

// Horizontal (uniform dx) matrices: compile-time constants folded into FMA immediates / constant bank.
template <int ORD> struct HMat {
  static constexpr bool is_const = true;
  __device__ __forceinline__ double lo(int i, int s, int ii) const { return Cst<ORD>::lo(i, s, ii); }
  __device__ __forceinline__ double br(int s, int ii) const { return Cst<ORD>::br(s, ii); }
};

// Vertical (variable dz) matrices: runtime table of one level, element stride `st`
// (1 when all ensemble members share one vertical grid, nens otherwise). Layout per level: lo[HS][HS][HS], br[ORD][ORD].
template <int ORD> struct VMat {
  static constexpr bool is_const = false;
  static constexpr int HS = (ORD + 1) / 2;
  static constexpr int NEL = HS * HS * HS + ORD * ORD;
  const double *p;
  long st;
  __device__ __forceinline__ double lo(int i, int s, int ii) const { return __ldg(p + ((i * HS + s) * HS + ii) * st); }
  __device__ __forceinline__ double br(int s, int ii) const { return __ldg(p + (HS * HS * HS + s * ORD + ii) * st); }
};

// Vertical matrices of a grid shared by all ensemble members, held in the constant bank: the level index is uniform
// across a warp of the marching kernels, so every element is one constant-cache load (no vector-memory traffic).
// Same per-level layout as VMat. Uploaded by OrdLaunch<ORD>::upload_cvtab.
#ifndef AWFL_CVTAB_DOUBLES
#define AWFL_CVTAB_DOUBLES 7680
#endif
static __constant__ double c_vtab[AWFL_CVTAB_DOUBLES];
template <int ORD> struct CMat {
  static constexpr bool is_const = false;
  static constexpr int HS = (ORD + 1) / 2;
  static constexpr int NEL = HS * HS * HS + ORD * ORD;
  int base;
  __device__ __forceinline__ double lo(int i, int s, int ii) const { return c_vtab[base + (i * HS + s) * HS + ii]; }
  __device__ __forceinline__ double br(int s, int ii) const { return c_vtab[base + HS * HS * HS + s * ORD + ii]; }
};

// WENO-limited polynomial of the centre cell of stencil u[0..ORD-1], returned as its even (E) and odd (O) parts at
// the cell edges: value(left edge, x=-1/2) = E - O ; value(right edge, x=+1/2) = E + O.
template <int ORD, class M>
__device__ __forceinline__ void weno_EO(const double (&u)[ORD], const M &m, double &E, double &O) {
  constexpr int HS = (ORD + 1) / 2;
  using C = Cst<ORD>;
  double alo[HS][HS], ab[ORD];
#pragma unroll
  for (int i = 0; i < HS; i++) {
#pragma unroll
    for (int ii = 0; ii < HS; ii++) {
      double t = m.lo(i, 0, ii) * u[i];
#pragma unroll
      for (int s = 1; s < HS; s++) {
        if (!M::is_const || m.lo(i, s, ii) != 0.0) t = fma(m.lo(i, s, ii), u[i + s], t);
      }
      alo[i][ii] = t;
    }
  }
#pragma unroll
  for (int ii = 0; ii < ORD; ii++) {
    double t = m.br(0, ii) * u[0];
#pragma unroll
    for (int s = 1; s < ORD; s++) {
      if (!M::is_const || m.br(s, ii) != 0.0) t = fma(m.br(s, ii), u[s], t);
    }
    ab[ii] = t;
  }
  double tv[HS + 1];
#pragma unroll
  for (int i = 0; i < HS; i++) tv[i] = C::tv_lo(alo[i]);
  tv[HS] = C::tv_hi(ab);
  double avg = tv[0];
#pragma unroll
  for (int i = 1; i < HS; i++) avg += tv[i];
  avg *= (1.0 / HS);
  tv[HS] = fma(tv[HS] - avg, C::sigma(), avg);
  double w[HS + 1], sum = 0.0;
#pragma unroll
  for (int i = 0; i < HS + 1; i++) {
    w[i] = C::idl(i) * rcp_pos(fma(tv[i], tv[i], 1.0e-20));
    sum += w[i];
  }
  const double r1 = rcp_pos(sum + 1.0e-20);
  double sum2 = 0.0;
#pragma unroll
  for (int i = 0; i < HS + 1; i++) {
    const double d = C::idl(i);
    const double x = w[i] * r1;
    const double num = x * fma(x, x - 3.0 * d, d + d * d);
    const double den = fma(x, 1.0 - 2.0 * d, d * d);
    w[i] = num * rcp_pos(den);
    sum2 += w[i];
  }
  const double r2 = rcp_pos(sum2 + 1.0e-20);
  double Eb = 0.0, Ob = 0.0;
#pragma unroll
  for (int i = 0; i < HS; i++) {
    double e = alo[i][0], o = 0.0;
    double sc = 0.5;
#pragma unroll
    for (int s = 1; s < HS; s++) {
      if (s & 1) o = fma(sc, alo[i][s], o); else e = fma(sc, alo[i][s], e);
      sc *= 0.5;
    }
    Eb = fma(w[i], e, Eb);
    Ob = fma(w[i], o, Ob);
  }
  {
    double e = ab[0], o = 0.0;
    double sc = 0.5;
#pragma unroll
    for (int s = 1; s < ORD; s++) {
      if (s & 1) o = fma(sc, ab[s], o); else e = fma(sc, ab[s], e);
      sc *= 0.5;
    }
    Eb = fma(w[HS], e, Eb);
    Ob = fma(w[HS], o, Ob);
  }
  E = Eb * r2;
  O = Ob * r2;
}

}  // namespace awfl
