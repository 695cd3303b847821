// FP64-pipe microbenchmarks for B200 (sm_100a): how does DFMA throughput behave when other instructions share the issue
// port, with distinct register operands, with constant-bank operands, at various occupancies?
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_mix fp64_mix.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

__constant__ double cc[64];

template <int MODE>
__global__ void __launch_bounds__(128) k(int iters, double seed, double *sink, int *isink) {
  double x0 = seed + threadIdx.x * 1e-9, x1 = x0 + 1e-3, x2 = x0 + 2e-3, x3 = x0 + 3e-3, x4 = x0 + 4e-3, x5 = x0 + 5e-3, x6 = x0 + 6e-3, x7 = x0 + 7e-3;
  const double a = 0.9999999, b = 1e-7;
  int i0 = threadIdx.x, i1 = i0 + 1, i2 = i0 + 2, i3 = i0 + 3;
  float f0 = threadIdx.x, f1 = f0 + 1.f;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      if (MODE == 0) {            // pure DFMA, shared operands
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
      } else if (MODE == 1) {     // DFMA, three distinct register operands
        x0 = fma(x1, x2, x0); x1 = fma(x2, x3, x1); x2 = fma(x3, x4, x2); x3 = fma(x4, x5, x3);
        x4 = fma(x5, x6, x4); x5 = fma(x6, x7, x5); x6 = fma(x7, x0, x6); x7 = fma(x0, x1, x7);
      } else if (MODE == 2) {     // DFMA : integer 1:1
        x0 = fma(x0, a, b); i0 = i0 * 3 + i1; x1 = fma(x1, a, b); i1 = i1 * 5 + i2; x2 = fma(x2, a, b); i2 = i2 * 7 + i3; x3 = fma(x3, a, b); i3 = i3 * 9 + i0;
        x4 = fma(x4, a, b); i0 = i0 ^ i2; x5 = fma(x5, a, b); i1 = i1 + i3; x6 = fma(x6, a, b); i2 = i2 ^ i0; x7 = fma(x7, a, b); i3 = i3 + i1;
      } else if (MODE == 3) {     // DFMA with constant-bank operands
        x0 = fma(x0, cc[0], cc[1]); x1 = fma(x1, cc[2], cc[3]); x2 = fma(x2, cc[4], cc[5]); x3 = fma(x3, cc[6], cc[7]);
        x4 = fma(x4, cc[8], cc[9]); x5 = fma(x5, cc[10], cc[11]); x6 = fma(x6, cc[12], cc[13]); x7 = fma(x7, cc[14], cc[15]);
      } else if (MODE == 4) {     // DMUL / DADD mix
        x0 = x0 * a; x1 = x1 + b; x2 = x2 * a; x3 = x3 + b; x4 = x4 * a; x5 = x5 + b; x6 = x6 * a; x7 = x7 + b;
      } else if (MODE == 5) {     // DFMA : FP32 FMA 1:1
        x0 = fma(x0, a, b); f0 = fmaf(f0, 0.999f, f1); x1 = fma(x1, a, b); f1 = fmaf(f1, 0.999f, f0); x2 = fma(x2, a, b); f0 = fmaf(f0, 0.99f, f1); x3 = fma(x3, a, b); f1 = fmaf(f1, 0.99f, f0);
        x4 = fma(x4, a, b); f0 = fmaf(f0, 0.999f, f1); x5 = fma(x5, a, b); f1 = fmaf(f1, 0.999f, f0); x6 = fma(x6, a, b); f0 = fmaf(f0, 0.99f, f1); x7 = fma(x7, a, b); f1 = fmaf(f1, 0.99f, f0);
      } else if (MODE == 6) {     // DFMA : integer 1:2
        x0 = fma(x0, a, b); i0 = i0 * 3 + i1; i1 = i1 * 5 + i2; x1 = fma(x1, a, b); i2 = i2 * 7 + i3; i3 = i3 * 9 + i0; x2 = fma(x2, a, b); i0 = i0 ^ i2; i1 = i1 + i3; x3 = fma(x3, a, b); i2 = i2 ^ i0; i3 = i3 + i1;
        x4 = fma(x4, a, b); i0 = i0 * 3 + i1; i1 = i1 * 5 + i2; x5 = fma(x5, a, b); i2 = i2 * 7 + i3; i3 = i3 * 9 + i0; x6 = fma(x6, a, b); i0 = i0 ^ i2; i1 = i1 + i3; x7 = fma(x7, a, b); i2 = i2 ^ i0; i3 = i3 + i1;
      } else if (MODE == 7) {     // single dependent chain per thread (latency)
        x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b);
      } else if (MODE == 8) {     // two chains
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x0 = fma(x0, a, b); x1 = fma(x1, a, b); x0 = fma(x0, a, b); x1 = fma(x1, a, b); x0 = fma(x0, a, b); x1 = fma(x1, a, b);
      }
    }
  }
  const double s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 12345.6789) sink[0] = s;
  if (i0 + i1 + i2 + i3 == 123456789 || f0 + f1 == 1.2345f) isink[0] = i0;
}

template <int MODE> double run(int blocks_per_sm, int threads, int iters, double *sink, int *isink) {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<sms * blocks_per_sm, threads>>>(iters / 8, 1.0, sink, isink);
  cudaEventRecord(e0);
  k<MODE><<<sms * blocks_per_sm, threads>>>(iters, 1.0, sink, isink);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double n = 64.0 * iters * (double)sms * blocks_per_sm * threads;   // FP64 instructions (thread-level)
  return n / (ms * 1e-3) / 1e12;   // T FP64-inst/s (x2 = TFLOP/s for FMA)
}

int main() {
  double *sink; int *isink; cudaMalloc(&sink, 8); cudaMalloc(&isink, 4);
  double h[64]; for (int i = 0; i < 64; i++) h[i] = i % 2 ? 1e-7 : 0.9999999; cudaMemcpyToSymbol(cc, h, sizeof(h));
  const char *names[] = {"pure DFMA shared operands", "DFMA 3 distinct reg operands", "DFMA:INT 1:1", "DFMA const-bank operands", "DMUL/DADD mix",
                         "DFMA:FFMA 1:1", "DFMA:INT 1:2", "1 chain/thread (latency)", "2 chains/thread"};
  const int iters = 4096;
  for (int warps_per_sm : {4, 8, 16, 32}) {
    const int bps = warps_per_sm / 4;
    printf("warps/SM=%d (T FP64 inst/s; peak ~18.6)\n", warps_per_sm);
    double v[9];
    v[0] = run<0>(bps, 128, iters, sink, isink); v[1] = run<1>(bps, 128, iters, sink, isink); v[2] = run<2>(bps, 128, iters, sink, isink);
    v[3] = run<3>(bps, 128, iters, sink, isink); v[4] = run<4>(bps, 128, iters, sink, isink); v[5] = run<5>(bps, 128, iters, sink, isink);
    v[6] = run<6>(bps, 128, iters, sink, isink); v[7] = run<7>(bps, 128, iters, sink, isink); v[8] = run<8>(bps, 128, iters, sink, isink);
    for (int m = 0; m < 9; m++) printf("  %-32s %.2f\n", names[m], v[m]);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
