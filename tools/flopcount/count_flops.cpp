// Instrumented operation count of one Dycore::reconstruct ("regenerate with a counting scalar type and freeze
// it in roofline.py"). The algorithm is pam_b200/cxx/Dycore.h's reconstruct_generic -- the reference's compute_weno_coefs +
// edge evaluation operation for operation (WenoLimiter.h:98-181, Dycore.h:599-603), bit-equal to the oracle (tests/
// test_cxx_shim.py) -- instantiated with a scalar type that counts every add/sub, mul and div it performs.
//   g++ -std=c++17 -DMW_ORD=5 -I pam_b200/cxx tools/flopcount/count_flops.cpp -o /tmp/count5 && /tmp/count5
#include <cstdio>
#include <cstring>
#include "Dycore.h"

struct Cnt {
  double v;
  static long add, mul, div;
  Cnt() : v(0) {}
  Cnt(double x) : v(x) {}
  friend Cnt operator+(Cnt a, Cnt b) { add++; return Cnt(a.v + b.v); }
  friend Cnt operator-(Cnt a, Cnt b) { add++; return Cnt(a.v - b.v); }
  friend Cnt operator*(Cnt a, Cnt b) { mul++; return Cnt(a.v * b.v); }
  friend Cnt operator/(Cnt a, Cnt b) { div++; return Cnt(a.v / b.v); }
};
long Cnt::add = 0, Cnt::mul = 0, Cnt::div = 0;

template <int D0, int D1 = 1, int D2 = 1> struct Mat {
  double v[D0 * D1 * D2];
  Cnt operator()(int i) const { return v[i]; }
  Cnt operator()(int i, int j) const { return v[i * D1 + j]; }
  Cnt operator()(int i, int j, int k) const { return v[(i * D1 + j) * D2 + k]; }
};
#define CAT3_(a, b, c) a##b##c
#define CAT3(a, b, c) CAT3_(a, b, c)

int main() {
  constexpr int ord = MW_ORD, hs = (ord + 1) / 2;
  Mat<ord, ord> s2c; Mat<ord, 2> c2g; Mat<hs, hs, hs> wrl; Mat<hs + 1> idl; Mat<ord> u;
  std::memcpy(s2c.v, CAT3(AWFL_S2C_, MW_ORD, ), sizeof(s2c.v));
  std::memcpy(c2g.v, CAT3(AWFL_C2G_, MW_ORD, ), sizeof(c2g.v));
  std::memcpy(wrl.v, CAT3(AWFL_WRL_, MW_ORD, ), sizeof(wrl.v));
  double sum = 0;
  for (int i = 0; i < hs + 1; i++) sum += CAT3(AWFL_IDL_RAW_, MW_ORD, )[i];
  for (int i = 0; i < hs + 1; i++) idl.v[i] = CAT3(AWFL_IDL_RAW_, MW_ORD, )[i] / (sum + 1.0e-20);
  for (int s = 0; s < ord; s++) u.v[s] = 1.0 + 0.1 * s * s;
  Cnt::add = Cnt::mul = Cnt::div = 0;
  Cnt r = awfl_b200_detail::reconstruct_generic<ord, Cnt>(u, c2g, s2c, wrl, idl, Cnt(CAT3(AWFL_SIGMA_, MW_ORD, )), 1);
  std::printf("{\"ord\": %d, \"add\": %ld, \"mul\": %ld, \"div\": %ld, \"flops\": %ld, \"value\": %.17g}\n", ord, Cnt::add, Cnt::mul, Cnt::div,
              Cnt::add + Cnt::mul + Cnt::div, r.v);
  return 0;
}
