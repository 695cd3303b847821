// FP64-pipe microbenchmarks for B200 (sm_100a), part 2: the operand forms the WENO dot products actually use, and the
// FP64 tensor-core instruction as an alternative.
//   MODE 0  acc_i = fma(x_i, a, acc_i)          a, b immediates shared by all (reference point, = fp64_mix mode 0)
//   MODE 1  acc_i = fma(x_i, cc[j], acc_i)      ONE constant-bank operand at a compile-time address + two registers
//   MODE 2  acc_i = fma(x_i, cc[base+j], acc_i) constant-bank operand at a uniform RUN-TIME index (LDCU + UR operand)
//   MODE 3  acc_i = fma(x_i, y_i, acc_i)        three registers, accumulator in place (two fresh + one reused)
//   MODE 4  acc_i = fma(x_i, y_j, z_k)          three registers, none shared with the destination
//   MODE 5  DMMA  mma.sync.m8n8k4.f64           (counted as 256 FMA per warp instruction = 8 per lane)
//   MODE 6  acc_i = fma(x_i, u, acc_i)          u in a uniform register, loaded ONCE outside the loop (pure UR-operand rate)
//   MODE 7  like 2, one coefficient load per TWO DFMA (paired reconstructions)   MODE 8: per FOUR
//   MODE 9  acc_i = fma(x_i, y, acc_i)          y a REGISTER shared by consecutive pairs (operand reuse cache)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_operands fp64_operands.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

__constant__ double cc[256];

template <int MODE>
__global__ void __launch_bounds__(128) k(int iters, int base, double seed, double *sink) {
  double a0 = seed, a1 = seed + 1, a2 = seed + 2, a3 = seed + 3, a4 = seed + 4, a5 = seed + 5, a6 = seed + 6, a7 = seed + 7;
  double x0 = 1e-9 * threadIdx.x, x1 = x0 + 1e-3, x2 = x0 + 2e-3, x3 = x0 + 3e-3, x4 = x0 + 4e-3, x5 = x0 + 5e-3, x6 = x0 + 6e-3, x7 = x0 + 7e-3;
  double y0 = x0 * 0.5, y1 = x1 * 0.5, y2 = x2 * 0.5, y3 = x3 * 0.5;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      if (MODE == 0) {
        a0 = fma(x0, 0.9999999, a0); a1 = fma(x1, 0.9999999, a1); a2 = fma(x2, 0.9999999, a2); a3 = fma(x3, 0.9999999, a3);
        a4 = fma(x4, 0.9999999, a4); a5 = fma(x5, 0.9999999, a5); a6 = fma(x6, 0.9999999, a6); a7 = fma(x7, 0.9999999, a7);
      } else if (MODE == 1) {
        a0 = fma(x0, cc[8 * r + 0], a0); a1 = fma(x1, cc[8 * r + 1], a1); a2 = fma(x2, cc[8 * r + 2], a2); a3 = fma(x3, cc[8 * r + 3], a3);
        a4 = fma(x4, cc[8 * r + 4], a4); a5 = fma(x5, cc[8 * r + 5], a5); a6 = fma(x6, cc[8 * r + 6], a6); a7 = fma(x7, cc[8 * r + 7], a7);
      } else if (MODE == 2) {
        const int b = base + 8 * r;
        a0 = fma(x0, cc[b + 0], a0); a1 = fma(x1, cc[b + 1], a1); a2 = fma(x2, cc[b + 2], a2); a3 = fma(x3, cc[b + 3], a3);
        a4 = fma(x4, cc[b + 4], a4); a5 = fma(x5, cc[b + 5], a5); a6 = fma(x6, cc[b + 6], a6); a7 = fma(x7, cc[b + 7], a7);
      } else if (MODE == 3) {
        a0 = fma(x0, y0, a0); a1 = fma(x1, y1, a1); a2 = fma(x2, y2, a2); a3 = fma(x3, y3, a3);
        a4 = fma(x4, y0, a4); a5 = fma(x5, y1, a5); a6 = fma(x6, y2, a6); a7 = fma(x7, y3, a7);
      } else if (MODE == 4) {
        a0 = fma(x1, y0, a4); a1 = fma(x2, y1, a5); a2 = fma(x3, y2, a6); a3 = fma(x4, y3, a7);
        a4 = fma(x5, y0, a0); a5 = fma(x6, y1, a1); a6 = fma(x7, y2, a2); a7 = fma(x0, y3, a3);
      } else if (MODE == 6) {
        const double u0 = cc[base], u1 = cc[base + 1];
        a0 = fma(x0, u0, a0); a1 = fma(x1, u1, a1); a2 = fma(x2, u0, a2); a3 = fma(x3, u1, a3);
        a4 = fma(x4, u0, a4); a5 = fma(x5, u1, a5); a6 = fma(x6, u0, a6); a7 = fma(x7, u1, a7);
      } else if (MODE == 7) {
        const int b = base + 4 * r;
        a0 = fma(x0, cc[b + 0], a0); a1 = fma(x1, cc[b + 0], a1); a2 = fma(x2, cc[b + 1], a2); a3 = fma(x3, cc[b + 1], a3);
        a4 = fma(x4, cc[b + 2], a4); a5 = fma(x5, cc[b + 2], a5); a6 = fma(x6, cc[b + 3], a6); a7 = fma(x7, cc[b + 3], a7);
      } else if (MODE == 8) {
        const int b = base + 2 * r;
        a0 = fma(x0, cc[b + 0], a0); a1 = fma(x1, cc[b + 0], a1); a2 = fma(x2, cc[b + 0], a2); a3 = fma(x3, cc[b + 0], a3);
        a4 = fma(x4, cc[b + 1], a4); a5 = fma(x5, cc[b + 1], a5); a6 = fma(x6, cc[b + 1], a6); a7 = fma(x7, cc[b + 1], a7);
      } else if (MODE == 9) {
        a0 = fma(x0, y0, a0); a1 = fma(x1, y0, a1); a2 = fma(x2, y1, a2); a3 = fma(x3, y1, a3);
        a4 = fma(x4, y2, a4); a5 = fma(x5, y2, a5); a6 = fma(x6, y3, a6); a7 = fma(x7, y3, a7);
      } else if (MODE == 5) {
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(a0), "+d"(a1) : "d"(x0), "d"(y0));
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(a2), "+d"(a3) : "d"(x1), "d"(y1));
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(a4), "+d"(a5) : "d"(x2), "d"(y2));
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(a6), "+d"(a7) : "d"(x3), "d"(y3));
      }
    }
    if (MODE == 2 || MODE == 7 || MODE == 8) base = (base + 64) & 127;
  }
  const double s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (s == 12345.6789) sink[0] = s;
}

template <int MODE> double run(int blocks_per_sm, int iters, double *sink) {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<sms * blocks_per_sm, 128>>>(iters / 8, 0, 1.0, sink);
  cudaEventRecord(e0);
  k<MODE><<<sms * blocks_per_sm, 128>>>(iters, 0, 1.0, sink);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double per_iter = MODE == 5 ? 8.0 * 4 * 8 : 64.0;   // FMA per thread and iteration (a DMMA m8n8k4 = 8 FMA per lane)
  return per_iter * iters * (double)sms * blocks_per_sm * 128 / (ms * 1e-3) / 1e12;   // T FMA/s (x2 = TFLOP/s)
}

int main() {
  double *sink; cudaMalloc(&sink, 8);
  double h[256]; for (int i = 0; i < 256; i++) h[i] = 1e-7 * (i + 1); cudaMemcpyToSymbol(cc, h, sizeof(h));
  const char *names[] = {"acc += x * imm", "acc += x * c[const addr]", "acc += x * c[uniform run-time index]", "acc += x * y (3 regs, acc in place)",
                         "d = x * y + z (3 regs, none reused)", "DMMA m8n8k4 (FMA-equivalents)", "acc += x * UR (loop-invariant UR)",
                         "acc += x * c[uniform idx], 1 load : 2 DFMA", "acc += x * c[uniform idx], 1 load : 4 DFMA", "acc += x * y, y shared by pairs (reuse)"};
  const int iters = 4096;
  for (int warps_per_sm : {4, 8, 12, 16}) {
    const int bps = warps_per_sm / 4;
    printf("warps/SM=%d (T FMA/s; DFMA peak ~18.4)\n", warps_per_sm);
    double v[10] = {run<0>(bps, iters, sink), run<1>(bps, iters, sink), run<2>(bps, iters, sink), run<3>(bps, iters, sink), run<4>(bps, iters, sink), run<5>(bps, iters, sink),
                    run<6>(bps, iters, sink), run<7>(bps, iters, sink), run<8>(bps, iters, sink), run<9>(bps, iters, sink)};
    for (int m = 0; m < 10; m++) printf("  %-40s %.2f\n", names[m], v[m]);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
