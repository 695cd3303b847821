// Device-side WENO reconstruction for the AWFL dycore (sm_100a).
//
// Mathematically identical to the reference's Dycore::reconstruct (dynamics/awfl/Dycore.h:591-604) +
// weno::compute_weno_coefs (dynamics/awfl/WenoLimiter.h:98-181), re-associated for the FP64 pipe:
//   * the "bridge" polynomial (WenoLimiter.h:131-138) is one mat-vec with a precomputed matrix
//     B = (sten_to_coefs - sum_i idl_i*embed(weno_lower_i)) / idl_m   (tools/gen_tables.py; per level on the host
//     for the vertical matrices), which removes ord divides and hs*hs FMAs per reconstruction;
//   * no candidate polynomial is formed coefficient by coefficient: its stencil is mapped straight to the even part of
//     its edge values and to the factors p_k of its total variation (TV = sum p_k^2, L D L^T of coefs_to_tv; the odd
//     part of the edge values is p_1/2) -- 131 FP64 instructions per ord-5 reconstruction on a uniform grid instead of
//     156 (the bridge rows are even / odd in the stencil there), 144 with per-level vertical matrices;
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

// Every table exists twice: a constexpr copy (compile-time structure: which entries are zero, which rows are even / odd)
// and a __constant__ copy whose entries the FP64 instructions read straight from the constant bank as c[bank][offset]
// operands. (Folded into the instruction stream as immediates, each 64-bit constant costs two extra MOVs per use:
// profiles/r01b measured as many non-FP64 as FP64 instructions in the x march.)
//
// "Row" form of a candidate polynomial (tools/gen_tables.py, tv_rows): ONE matrix applied to the stencil gives
//   row 0      e   = even part of the edge values            p(-1/2) = e - o,  p(+1/2) = e + o
//   rows 1..   p_k = the factors of the truncated total variation, TV = sum_k p_k^2   (L D L^T of coefs_to_tv)
// and the odd part is o = p_1 / 2 for every order, so the polynomial coefficients themselves are never formed:
// rlo[i][r][s] for the (ord+1)/2 low-order stencils, rhi[r][s] for the bridge polynomial (WenoLimiter.h:131-138 folded
// into the matrix). ghi/glo are the same maps in coefficient space, applied on the host to the per-level vertical
// matrices.
#define AWFL_DEFINE_CST(O)                                                                                      \
  namespace ctab##O {                                                                                           \
    constexpr int HS = (O + 1) / 2;                                                                             \
    constexpr double idl_h[HS + 1] = AWFL_IDL_##O##_INIT;                                                       \
    static __constant__ double rlo[HS][HS][HS] = AWFL_RLO_##O##_INIT;                                           \
    static __constant__ double rhi[O][O] = AWFL_RHI_##O##_INIT;                                                 \
    static __constant__ double idl[HS + 1] = AWFL_IDL_##O##_INIT;                                               \
    static __constant__ MapTab<HS + 1> map = make_map<HS + 1>(idl_h);                                           \
    static __constant__ double misc[3] = {AWFL_SIGMA_##O, 1.0 / HS, 1.0e-20};                                   \
  }                                                                                                             \
  template <> struct Cst<O> {                                                                                   \
    static constexpr int HS = (O + 1) / 2;                                                                      \
    static __device__ __forceinline__ bool lo_nz(int i, int r, int s) {                                         \
      constexpr double T[HS][HS][HS] = AWFL_RLO_##O##_INIT;                                                     \
      return T[i][r][s] != 0.0;                                                                                 \
    }                                                                                                           \
    static __device__ __forceinline__ bool hi_nz(int r, int s) {                                                \
      constexpr double T[O][O] = AWFL_RHI_##O##_INIT;                                                           \
      return T[r][s] != 0.0;                                                                                    \
    }                                                                                                           \
    static __device__ __forceinline__ bool hi_even(int r) {                                                     \
      constexpr int T[O] = AWFL_RHI_PAR_##O##_INIT;                                                             \
      return T[r] > 0;                                                                                          \
    }                                                                                                           \
    static __device__ __forceinline__ double lo(int i, int r, int s) { return ctab##O::rlo[i][r][s]; }          \
    static __device__ __forceinline__ double hi(int r, int s) { return ctab##O::rhi[r][s]; }                    \
    static __device__ __forceinline__ double idl(int i) { return ctab##O::idl[i]; }                             \
    static __device__ __forceinline__ double map(int i, int k) { return ctab##O::map.v[i][k]; }                 \
    static __device__ __forceinline__ double sigma() { return ctab##O::misc[0]; }                               \
    static __device__ __forceinline__ double rhs() { return ctab##O::misc[1]; }                                 \
    static __device__ __forceinline__ double eps() { return ctab##O::misc[2]; }                                 \
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

// Horizontal (uniform dx) matrices: constant-bank tables with compile-time structure (zeros, even / odd bridge rows).
template <int ORD> struct HMat {
  static constexpr bool is_const = true;
  __device__ __forceinline__ double lo(int i, int r, int s) const { return Cst<ORD>::lo(i, r, s); }
  __device__ __forceinline__ double hi(int r, int s) const { return Cst<ORD>::hi(r, s); }
  __device__ __forceinline__ bool lo_nz(int i, int r, int s) const { return Cst<ORD>::lo_nz(i, r, s); }
  __device__ __forceinline__ bool hi_nz(int r, int s) const { return Cst<ORD>::hi_nz(r, s); }
};

// Vertical (variable dz) matrices: runtime table of one level, element stride `st`
// (1 when all ensemble members share one vertical grid, nens otherwise). Layout per level: rlo[HS][HS][HS], rhi[ORD][ORD].
template <int ORD> struct VMat {
  static constexpr bool is_const = false;
  static constexpr int HS = (ORD + 1) / 2;
  static constexpr int NEL = HS * HS * HS + ORD * ORD;
  const double *p;
  long st;
  __device__ __forceinline__ double lo(int i, int r, int s) const { return __ldg(p + ((i * HS + r) * HS + s) * st); }
  __device__ __forceinline__ double hi(int r, int s) const { return __ldg(p + (HS * HS * HS + r * ORD + s) * st); }
};

// The same table for a grid shared by all ensemble members when it does not fit the constant bank (ord 9 at 58 levels):
// unit element stride known at compile time, so every element is a load at an immediate offset from the level's base --
// no 64-bit address arithmetic per element (with a run-time stride the ord-9 z march executed more IMAD than DFMA).
template <int ORD> struct UMat {
  static constexpr bool is_const = false;
  static constexpr int HS = (ORD + 1) / 2;
  static constexpr int NEL = HS * HS * HS + ORD * ORD;
  const double *p;
  __device__ __forceinline__ double lo(int i, int r, int s) const { return __ldg(p + ((i * HS + r) * HS + s)); }
  __device__ __forceinline__ double hi(int r, int s) const { return __ldg(p + (HS * HS * HS + r * ORD + s)); }
};

// Vertical matrices of a grid shared by all ensemble members, held in the constant bank: the level index is uniform
// across a warp of the marching kernels, so every element is one constant-cache load (no vector-memory traffic).
// Same per-level layout as VMat. Uploaded by OrdLaunch<ORD>::upload_cvtab.
// Levels whose tables are bit-identical share one copy (on an equally spaced grid that is every level, and then even the
// 206-double ord-9 table of 60 levels fits): c_vlvl[level] is the offset of the level's table in c_vtab.
#ifndef AWFL_CVTAB_DOUBLES
#define AWFL_CVTAB_DOUBLES 7168
#endif
#ifndef AWFL_CVLVL_MAX
#define AWFL_CVLVL_MAX 512
#endif
static __constant__ double c_vtab[AWFL_CVTAB_DOUBLES];
static __constant__ int c_vlvl[AWFL_CVLVL_MAX];
template <int ORD> struct CMat {
  static constexpr bool is_const = false;
  static constexpr int HS = (ORD + 1) / 2;
  static constexpr int NEL = HS * HS * HS + ORD * ORD;
  int base;
  __device__ __forceinline__ double lo(int i, int r, int s) const { return c_vtab[base + (i * HS + r) * HS + s]; }
  __device__ __forceinline__ double hi(int r, int s) const { return c_vtab[base + HS * HS * HS + r * ORD + s]; }
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

// sum of N values as a balanced tree (dependency depth log2 N instead of N: the weight computation is the
// latency-critical part of a reconstruction)
template <int N> __device__ __forceinline__ double tree_sum(const double (&x)[N]) {
  double y[N];
#pragma unroll
  for (int i = 0; i < N; i++) y[i] = x[i];
#pragma unroll
  for (int w = 1; w < N; w *= 2) {
#pragma unroll
    for (int i = 0; i + w < N; i += 2 * w) y[i] += y[i + w];
  }
  return y[0];
}

template <class M> __device__ __forceinline__ bool nonzero_lo(const M &m, int i, int r, int s) {
  if constexpr (M::is_const) return m.lo_nz(i, r, s); else return true;
}
template <class M> __device__ __forceinline__ bool nonzero_hi(const M &m, int r, int s) {
  if constexpr (M::is_const) return m.hi_nz(r, s); else return true;
}

// WENO-limited polynomials of the centre cell of NV stencils u[v][0..ORD-1] at once, each returned as its even (E) and odd
// (O) parts at the cell edges: value(left edge, x=-1/2) = E - O ; value(right edge, x=+1/2) = E + O. The NV
// reconstructions are independent; evaluating them together shares every table load between them (one uniform load per
// NV FP64 instructions) and gives the FP64 pipe NV independent dependency chains.
//
// SAFE = true evaluates the weights the way the reference writes them -- every weight divided on its own, no products of
// the (tv^2 + eps) factors -- which stays finite for data of any magnitude. The fast form multiplies hs+1 factors of order
// data^4 and overflows FP64 beyond |data| ~ 1e19 (ord 5) / 7e12 (ord 9); it then returns NaN (never a wrong finite value:
// any overflowing product poisons the common denominator), which the tracer march tests for and redoes with SAFE = true.
// The state march reconstructs rho*u, p, u, v, w, theta, whose physical magnitudes (<= 1e6) are far from that range.
template <int ORD, int NV, class M, bool SAFE = false>
__device__ __forceinline__ void weno_EO_n(const double (&u)[NV][ORD], const M &m, double (&E)[NV], double (&O)[NV]) {
  constexpr int HS = (ORD + 1) / 2;
  constexpr int N = HS + 1;
  using C = Cst<ORD>;
  // per candidate polynomial: even edge part e, twice the odd edge part o2 (= p_1), total variation tv
  double e[NV][N], o2[NV][N], tv[NV][N];
#pragma unroll
  for (int i = 0; i < HS; i++) {
#pragma unroll
    for (int r = 0; r < HS; r++) {
      double t[NV];
      bool started = false;
#pragma unroll
      for (int s = 0; s < HS; s++) {
        if (!nonzero_lo(m, i, r, s)) continue;
#pragma unroll
        for (int v = 0; v < NV; v++) t[v] = started ? fma(m.lo(i, r, s), u[v][i + s], t[v]) : m.lo(i, r, s) * u[v][i + s];
        started = true;
      }
#pragma unroll
      for (int v = 0; v < NV; v++) {
        if (r == 0) e[v][i] = t[v];
        else if (r == 1) { o2[v][i] = t[v]; tv[v][i] = t[v] * t[v]; }
        else tv[v][i] = fma(t[v], t[v], tv[v][i]);
      }
    }
  }
  // bridge polynomial. On a uniform grid its rows are even or odd in the stencil (RHI_PAR): they act on the sums /
  // differences of mirrored cells, HS resp. HS-1 terms instead of ORD.
  double sm[NV][HS], df[NV][HS];
  if constexpr (M::is_const) {
#pragma unroll
    for (int v = 0; v < NV; v++) {
#pragma unroll
      for (int k = 0; k < HS - 1; k++) { sm[v][k] = u[v][ORD - 1 - k] + u[v][k]; df[v][k] = u[v][ORD - 1 - k] - u[v][k]; }
      sm[v][HS - 1] = u[v][HS - 1];
    }
  }
#pragma unroll
  for (int r = 0; r < ORD; r++) {
    double t[NV];
    bool started = false;
    if constexpr (M::is_const) {
      const bool even = C::hi_even(r);
#pragma unroll
      for (int k = 0; k < HS; k++) {
        if (!even && k == HS - 1) continue;
        if (!nonzero_hi(m, r, ORD - 1 - k)) continue;
#pragma unroll
        for (int v = 0; v < NV; v++) {
          const double x = even ? sm[v][k] : df[v][k];
          t[v] = started ? fma(m.hi(r, ORD - 1 - k), x, t[v]) : m.hi(r, ORD - 1 - k) * x;
        }
        started = true;
      }
    } else {
#pragma unroll
      for (int s = 0; s < ORD; s++) {
#pragma unroll
        for (int v = 0; v < NV; v++) t[v] = started ? fma(m.hi(r, s), u[v][s], t[v]) : m.hi(r, s) * u[v][s];
        started = true;
      }
    }
#pragma unroll
    for (int v = 0; v < NV; v++) {
      if (r == 0) e[v][HS] = t[v];
      else if (r == 1) { o2[v][HS] = t[v]; tv[v][HS] = t[v] * t[v]; }
      else tv[v][HS] = fma(t[v], t[v], tv[v][HS]);
    }
  }
  const double eps = C::eps();
#pragma unroll
  for (int v = 0; v < NV; v++) {
    double tl[HS];
#pragma unroll
    for (int i = 0; i < HS; i++) tl[i] = tv[v][i];
    const double avg = tree_sum<HS>(tl) * C::rhs();
    tv[v][HS] = fma(tv[v][HS] - avg, C::sigma(), avg);
    if constexpr (SAFE) {
      double w[N], sum = 0.0;
#pragma unroll
      for (int i = 0; i < N; i++) { w[i] = C::idl(i) / fma(tv[v][i], tv[v][i], eps); sum += w[i]; }
      double sum2 = 0.0;
#pragma unroll
      for (int i = 0; i < N; i++) {
        const double d = C::idl(i), x = w[i] / (sum + eps);
        w[i] = x * (d + d * d - 3.0 * d * x + x * x) / (d * d + x * (1.0 - 2.0 * d));
        sum2 += w[i];
      }
      double Eb = 0.0, Ob = 0.0;
#pragma unroll
      for (int i = 0; i < N; i++) { const double x = w[i] / (sum2 + eps); Eb = fma(x, e[v][i], Eb); Ob = fma(x, o2[v][i], Ob); }
      E[v] = Eb;
      O[v] = 0.5 * Ob;
      continue;
    }
    // Weights (WenoLimiter.h:142-166): w_i = idl_i/(tv_i^2+eps); convexify; map_weights; convexify. Both normalisations
    // are carried out over a common denominator -- w_i/(sum_j w_j + eps) = idl_i P_i / (sum_j idl_j P_j + eps P) with
    // P_i = prod_{j!=i}(tv_j^2+eps), P = prod_j(tv_j^2+eps) -- the same rational expression evaluated with two
    // reciprocals instead of 2(hs+1)+2 (every factor is >= eps resp. >= idl^2, and <= ~1e40 for |data| <= 1e10: the
    // products neither overflow nor underflow in FP64).
    double t[N], pr[N], x[N];
#pragma unroll
    for (int i = 0; i < N; i++) t[i] = fma(tv[v][i], tv[v][i], eps);
    others_product<N>(t, pr);
#pragma unroll
    for (int i = 0; i < N; i++) x[i] = C::idl(i) * pr[i];
    const double r1 = rcp_pos(fma(eps, pr[0] * t[0], tree_sum<N>(x)));
    double num[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
      // map_weights (WenoLimiter.h:60-70): x (d + d^2 - 3 d x + x^2) / (d^2 + x (1 - 2 d))
      const double xi = x[i] * r1;
      num[i] = xi * fma(xi, xi - C::map(i, 0), C::map(i, 1));
      t[i] = fma(xi, C::map(i, 2), C::map(i, 3));
    }
    others_product<N>(t, pr);
    double w[N];
#pragma unroll
    for (int i = 0; i < N; i++) w[i] = num[i] * pr[i];
    const double r2 = rcp_pos(fma(eps, pr[0] * t[0], tree_sum<N>(w)));
    double Eb = w[0] * e[v][0], Ob = w[0] * o2[v][0];
#pragma unroll
    for (int i = 1; i < N; i++) { Eb = fma(w[i], e[v][i], Eb); Ob = fma(w[i], o2[v][i], Ob); }
    E[v] = Eb * r2;
    O[v] = Ob * (0.5 * r2);
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
