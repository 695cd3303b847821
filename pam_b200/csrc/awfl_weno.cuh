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

// map_weights constants per ideal weight d (WenoLimiter.h:60-70): {3d, d+d^2, 1-2d, d^2}
template <int N> struct MapTab { double v[N][4]; };
template <int N> constexpr MapTab<N> make_map(const double (&d)[N]) {
  MapTab<N> t{};
  for (int i = 0; i < N; i++) {
    t.v[i][0] = 3.0 * d[i]; t.v[i][1] = d[i] + d[i] * d[i]; t.v[i][2] = 1.0 - 2.0 * d[i]; t.v[i][3] = d[i] * d[i];
  }
  return t;
}

// Every table exists twice: a constexpr copy (compile-time structure: which entries are zero) and a __constant__ copy
// whose entries the FP64 instructions read straight from the constant bank as c[bank][offset] operands. (Folded into
// the instruction stream as immediates, each 64-bit constant costs two extra MOVs per use: profiles/r01b measured as
// many non-FP64 as FP64 instructions in the x march.)
#define AWFL_DEFINE_CST(O)                                                                                      \
  namespace ctab##O {                                                                                           \
    constexpr int HS = (O + 1) / 2;                                                                             \
    constexpr double idl_h[HS + 1] = AWFL_IDL_##O##_INIT;                                                       \
    static __constant__ double lo[HS][HS][HS] = AWFL_WRL_##O##_INIT;                                            \
    static __constant__ double br[O][O] = AWFL_BRIDGE_##O##_INIT;                                               \
    static __constant__ double idl[HS + 1] = AWFL_IDL_##O##_INIT;                                               \
    static __constant__ MapTab<HS + 1> map = make_map<HS + 1>(idl_h);                                           \
    static __constant__ double tvlo[AWFL_TVLO_N_##O] = AWFL_TVLO_C_##O##_INIT;                                  \
    static __constant__ double tvhi[AWFL_TVHI_N_##O] = AWFL_TVHI_C_##O##_INIT;                                  \
    static __constant__ double misc[3] = {AWFL_SIGMA_##O, 1.0 / HS, 1.0e-20};                                   \
  }                                                                                                             \
  template <> struct Cst<O> {                                                                                   \
    static constexpr int HS = (O + 1) / 2;                                                                      \
    static __device__ __forceinline__ bool lo_nz(int i, int s, int ii) {                                        \
      constexpr double T[HS][HS][HS] = AWFL_WRL_##O##_INIT;                                                     \
      return T[i][s][ii] != 0.0;                                                                                \
    }                                                                                                           \
    static __device__ __forceinline__ bool br_nz(int s, int ii) {                                               \
      constexpr double T[O][O] = AWFL_BRIDGE_##O##_INIT;                                                        \
      return T[s][ii] != 0.0;                                                                                   \
    }                                                                                                           \
    static __device__ __forceinline__ double lo(int i, int s, int ii) { return ctab##O::lo[i][s][ii]; }         \
    static __device__ __forceinline__ double br(int s, int ii) { return ctab##O::br[s][ii]; }                   \
    static __device__ __forceinline__ double idl(int i) { return ctab##O::idl[i]; }                             \
    static __device__ __forceinline__ double map(int i, int k) { return ctab##O::map.v[i][k]; }                 \
    static __device__ __forceinline__ double sigma() { return ctab##O::misc[0]; }                               \
    static __device__ __forceinline__ double rhs() { return ctab##O::misc[1]; }                                 \
    static __device__ __forceinline__ double eps() { return ctab##O::misc[2]; }                                 \
    /* TransformMatrices.h coefs_to_tv for hs and ord coefficients (truncated squared-derivative integrals) */  \
    static __device__ __forceinline__ double tv_lo(const double (&a)[HS]) {                                     \
      constexpr int N = AWFL_TVLO_N_##O;                                                                        \
      constexpr int I[N] = AWFL_TVLO_I_##O##_INIT;                                                              \
      constexpr int J[N] = AWFL_TVLO_J_##O##_INIT;                                                              \
      constexpr double C[N] = AWFL_TVLO_C_##O##_INIT;                                                           \
      double r = C[0] == 1.0 ? a[I[0]] * a[J[0]] : (ctab##O::tvlo[0] * a[I[0]]) * a[J[0]];                      \
      _Pragma("unroll") for (int t = 1; t < N; t++) r = fma(ctab##O::tvlo[t] * a[I[t]], a[J[t]], r);            \
      return r;                                                                                                 \
    }                                                                                                           \
    static __device__ __forceinline__ double tv_hi(const double (&a)[O]) {                                      \
      constexpr int N = AWFL_TVHI_N_##O;                                                                        \
      constexpr int I[N] = AWFL_TVHI_I_##O##_INIT;                                                              \
      constexpr int J[N] = AWFL_TVHI_J_##O##_INIT;                                                              \
      constexpr double C[N] = AWFL_TVHI_C_##O##_INIT;                                                           \
      double r = C[0] == 1.0 ? a[I[0]] * a[J[0]] : (ctab##O::tvhi[0] * a[I[0]]) * a[J[0]];                      \
      _Pragma("unroll") for (int t = 1; t < N; t++) r = fma(ctab##O::tvhi[t] * a[I[t]], a[J[t]], r);            \
      return r;                                                                                                 \
    }                                                                                                           \
  };

// one translation unit per order: only that order's tables take constant-bank space
#ifdef AWFL_ORD_INST
#if AWFL_ORD_INST == 3
AWFL_DEFINE_CST(3)
#elif AWFL_ORD_INST == 5
AWFL_DEFINE_CST(5)
#elif AWFL_ORD_INST == 7
AWFL_DEFINE_CST(7)
#else
AWFL_DEFINE_CST(9)
#endif
#else
AWFL_DEFINE_CST(3)
AWFL_DEFINE_CST(5)
AWFL_DEFINE_CST(7)
AWFL_DEFINE_CST(9)
#endif

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
  __device__ __forceinline__ bool lo_nz(int i, int s, int ii) const { return Cst<ORD>::lo_nz(i, s, ii); }
  __device__ __forceinline__ bool br_nz(int s, int ii) const { return Cst<ORD>::br_nz(s, ii); }
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
#define AWFL_CVTAB_DOUBLES 7424
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

// pr[i] = product of all t[j], j != i (prefix/suffix products, 3N-5 multiplications for N >= 3)
template <int N> __device__ __forceinline__ void others_product(const double (&t)[N], double (&pr)[N]) {
  double pre[N], suf[N];
  pre[0] = t[0];
#pragma unroll
  for (int i = 1; i < N - 1; i++) pre[i] = pre[i - 1] * t[i];
  suf[N - 1] = t[N - 1];
#pragma unroll
  for (int i = N - 2; i >= 1; i--) suf[i] = suf[i + 1] * t[i];
  pr[0] = suf[1];
  pr[N - 1] = pre[N - 2];
#pragma unroll
  for (int i = 1; i < N - 1; i++) pr[i] = pre[i - 1] * suf[i + 1];
}

template <class M> __device__ __forceinline__ bool nonzero_lo(const M &m, int i, int s, int ii) {
  if constexpr (M::is_const) return m.lo_nz(i, s, ii); else return true;
}
template <class M> __device__ __forceinline__ bool nonzero_br(const M &m, int s, int ii) {
  if constexpr (M::is_const) return m.br_nz(s, ii); else return true;
}

// WENO-limited polynomials of the centre cell of NV stencils u[v][0..ORD-1] at once, each returned as its even (E) and odd
// (O) parts at the cell edges: value(left edge, x=-1/2) = E - O ; value(right edge, x=+1/2) = E + O. The NV
// reconstructions are independent; evaluating them together shares every table load between them (one uniform load per
// NV FP64 instructions) and gives the FP64 pipe NV independent dependency chains.
template <int ORD, int NV, class M>
__device__ __forceinline__ void weno_EO_n(const double (&u)[NV][ORD], const M &m, double (&E)[NV], double (&O)[NV]) {
  constexpr int HS = (ORD + 1) / 2;
  constexpr int N = HS + 1;
  using C = Cst<ORD>;
  double alo[NV][HS][HS], ab[NV][ORD];
#pragma unroll
  for (int i = 0; i < HS; i++) {
#pragma unroll
    for (int ii = 0; ii < HS; ii++) {
      double t[NV];
#pragma unroll
      for (int v = 0; v < NV; v++) t[v] = m.lo(i, 0, ii) * u[v][i];
#pragma unroll
      for (int s = 1; s < HS; s++) {
        if (nonzero_lo(m, i, s, ii)) {
#pragma unroll
          for (int v = 0; v < NV; v++) t[v] = fma(m.lo(i, s, ii), u[v][i + s], t[v]);
        }
      }
#pragma unroll
      for (int v = 0; v < NV; v++) alo[v][i][ii] = t[v];
    }
  }
#pragma unroll
  for (int ii = 0; ii < ORD; ii++) {
    double t[NV];
#pragma unroll
    for (int v = 0; v < NV; v++) t[v] = m.br(0, ii) * u[v][0];
#pragma unroll
    for (int s = 1; s < ORD; s++) {
      if (nonzero_br(m, s, ii)) {
#pragma unroll
        for (int v = 0; v < NV; v++) t[v] = fma(m.br(s, ii), u[v][s], t[v]);
      }
    }
#pragma unroll
    for (int v = 0; v < NV; v++) ab[v][ii] = t[v];
  }
  const double eps = C::eps();
#pragma unroll
  for (int v = 0; v < NV; v++) {
    double tv[N];
#pragma unroll
    for (int i = 0; i < HS; i++) tv[i] = C::tv_lo(alo[v][i]);
    tv[HS] = C::tv_hi(ab[v]);
    double avg = tv[0];
#pragma unroll
    for (int i = 1; i < HS; i++) avg += tv[i];
    avg *= C::rhs();
    tv[HS] = fma(tv[HS] - avg, C::sigma(), avg);
    // Weights (WenoLimiter.h:142-166): w_i = idl_i/(tv_i^2+eps); convexify; map_weights; convexify. Both normalisations
    // are carried out over a common denominator -- w_i/(sum_j w_j + eps) = idl_i P_i / (sum_j idl_j P_j + eps P) with
    // P_i = prod_{j!=i}(tv_j^2+eps), P = prod_j(tv_j^2+eps) -- the same rational expression evaluated with two
    // reciprocals instead of 2(hs+1)+2 (every factor is >= eps resp. >= idl^2, and <= ~1e40 for |data| <= 1e10: the
    // products neither overflow nor underflow in FP64).
    double t[N], pr[N], x[N];
#pragma unroll
    for (int i = 0; i < N; i++) t[i] = fma(tv[i], tv[i], eps);
    others_product<N>(t, pr);
    double sum = eps * (pr[0] * t[0]);
#pragma unroll
    for (int i = 0; i < N; i++) { x[i] = C::idl(i) * pr[i]; sum += x[i]; }
    const double r1 = rcp_pos(sum);
    double num[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
      // map_weights (WenoLimiter.h:60-70): x (d + d^2 - 3 d x + x^2) / (d^2 + x (1 - 2 d))
      const double xi = x[i] * r1;
      num[i] = xi * fma(xi, xi - C::map(i, 0), C::map(i, 1));
      t[i] = fma(xi, C::map(i, 2), C::map(i, 3));
    }
    others_product<N>(t, pr);
    double w[N], sum2 = eps * (pr[0] * t[0]);
#pragma unroll
    for (int i = 0; i < N; i++) { w[i] = num[i] * pr[i]; sum2 += w[i]; }
    const double r2 = rcp_pos(sum2);
    double Eb = 0.0, Ob = 0.0;
#pragma unroll
    for (int i = 0; i < HS; i++) {
      double e = alo[v][i][0], o = 0.0;
      double sc = 0.5;
#pragma unroll
      for (int s = 1; s < HS; s++) {
        if (s & 1) o = fma(sc, alo[v][i][s], o); else e = fma(sc, alo[v][i][s], e);
        sc *= 0.5;
      }
      Eb = fma(w[i], e, Eb);
      Ob = fma(w[i], o, Ob);
    }
    {
      double e = ab[v][0], o = 0.0;
      double sc = 0.5;
#pragma unroll
      for (int s = 1; s < ORD; s++) {
        if (s & 1) o = fma(sc, ab[v][s], o); else e = fma(sc, ab[v][s], e);
        sc *= 0.5;
      }
      Eb = fma(w[HS], e, Eb);
      Ob = fma(w[HS], o, Ob);
    }
    E[v] = Eb * r2;
    O[v] = Ob * r2;
  }
}

template <int ORD, class M>
__device__ __forceinline__ void weno_EO(const double (&u)[ORD], const M &m, double &E, double &O) {
  double uu[1][ORD], EE[1], OO[1];
#pragma unroll
  for (int s = 0; s < ORD; s++) uu[0][s] = u[s];
  weno_EO_n<ORD, 1>(uu, m, EE, OO);
  E = EE[0]; O = OO[0];
}

}  // namespace awfl
