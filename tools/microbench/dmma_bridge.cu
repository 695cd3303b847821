// DMMA (mma.sync.m8n8k4.f64) prototype of the ord-9 "bridge" mat-vec of a WENO reconstruction, MEASURED against the
// FP64-pipe forms the marching kernels use ("measure rather than estimate").
//
// The linear map is y = B u, B 9x9 (row r of the row form: stencil -> (e, p_1 .. p_8)), one stencil per ensemble member,
// lane = member. Per warp and per set of 32 stencils:
//   mode 0  scalar, full rows          : 81 DFMA per lane, matrix entries as constant-bank operands
//   mode 1  scalar, even/odd rows      : the rows are even or odd in the stencil on a uniform grid: 8 adds for the mirrored sums
//                                        and differences, then 5 resp. 4 terms per row -> 41 DFMA/DMUL + 8 DADD per lane
//   mode 2  DMMA, operands in place    : Y(16 x 32) = B(16 x 12, zero-padded) U(12 x 32): 2 m-tiles x 4 n-tiles x 3 k-steps = 24
//                                        mma.sync.m8n8k4 per warp, the U fragments ALREADY in fragment layout (the best case:
//                                        no cost charged for bringing lane = member data into it, nor for bringing Y back)
//   mode 3  DMMA + layout exchange     : the same, U read from a [tap][member] shared-memory tile into fragment layout and Y
//                                        written back through shared memory to lane = member (what a marching kernel would need)
// Reported: stencil sets per second per GPU and the time relative to mode 1 (the form in use).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dmma_bridge dmma_bridge.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

__constant__ double cB[9][9];

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int MODE>
__global__ void __launch_bounds__(128) k(int iters, const double *__restrict__ uin, double *sink) {
  __shared__ double tile[4][12 * 32 + 16 * 32];        // per warp: U tile [12][32] (+ Y tile [16][32])
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double u[9];
#pragma unroll
  for (int s = 0; s < 9; s++) u[s] = uin[(blockIdx.x * blockDim.x + threadIdx.x) * 9 + s];
  double acc = 0.0;
  if (MODE == 0) {
    for (int it = 0; it < iters; it++) {
      double y[9];
#pragma unroll
      for (int r = 0; r < 9; r++) {
        double t = cB[r][0] * u[0];
#pragma unroll
        for (int s = 1; s < 9; s++) t = fma(cB[r][s], u[s], t);
        y[r] = t;
      }
#pragma unroll
      for (int r = 0; r < 9; r++) { acc += y[r]; u[r] = fma(y[r], 1e-9, u[r]); }
    }
  } else if (MODE == 1) {
    for (int it = 0; it < iters; it++) {
      double sm[5], df[4], y[9];
#pragma unroll
      for (int q = 0; q < 4; q++) { sm[q] = u[8 - q] + u[q]; df[q] = u[8 - q] - u[q]; }
      sm[4] = u[4];
#pragma unroll
      for (int r = 0; r < 9; r++) {
        double t;
        if (r % 2 == 0) {
          t = cB[r][0] * sm[0];
#pragma unroll
          for (int q = 1; q < 5; q++) t = fma(cB[r][q], sm[q], t);
        } else {
          t = cB[r][0] * df[0];
#pragma unroll
          for (int q = 1; q < 4; q++) t = fma(cB[r][q], df[q], t);
        }
        y[r] = t;
      }
#pragma unroll
      for (int r = 0; r < 9; r++) { acc += y[r]; u[r] = fma(y[r], 1e-9, u[r]); }
    }
  } else {
    // A fragments of B (zero-padded to 16 x 12): lane holds A[row = lane/4][col = lane%4] of each (m-tile, k-step)
    double a[2][3];
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
      for (int ks = 0; ks < 3; ks++) {
        const int r = mt * 8 + lane / 4, c = ks * 4 + lane % 4;
        a[mt][ks] = (r < 9 && c < 9) ? cB[r][c] : 0.0;
      }
    double *U = tile[w], *Y = tile[w] + 12 * 32;
    // B fragments of U: lane holds U[k = ks*4 + lane%4][member = nt*8 + lane/4]
    double b[4][3];
    if (MODE == 2) {
#pragma unroll
      for (int nt = 0; nt < 4; nt++)
#pragma unroll
        for (int ks = 0; ks < 3; ks++) b[nt][ks] = u[(nt + ks) % 9] + 1e-3 * (lane % 4);
    } else {
#pragma unroll
      for (int s = 9; s < 12; s++) U[s * 32 + lane] = 0.0;
    }
    for (int it = 0; it < iters; it++) {
      if (MODE == 3) {
#pragma unroll
        for (int s = 0; s < 9; s++) U[s * 32 + lane] = u[s];          // lane = member layout into the tile
        __syncwarp();
#pragma unroll
        for (int nt = 0; nt < 4; nt++)
#pragma unroll
          for (int ks = 0; ks < 3; ks++) b[nt][ks] = U[(ks * 4 + lane % 4) * 32 + nt * 8 + lane / 4];
      }
      double c[2][4][2];
#pragma unroll
      for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int nt = 0; nt < 4; nt++) {
          c[mt][nt][0] = 0.0; c[mt][nt][1] = 0.0;
#pragma unroll
          for (int ks = 0; ks < 3; ks++) dmma(c[mt][nt][0], c[mt][nt][1], a[mt][ks], b[nt][ks]);
        }
      if (MODE == 2) {
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
          for (int nt = 0; nt < 4; nt++) { acc += c[mt][nt][0] + c[mt][nt][1]; b[nt][mt] = fma(c[mt][nt][0], 1e-9, b[nt][mt]); }
      } else {
        // C fragment: row = mt*8 + lane/4, members nt*8 + 2*(lane%4) + {0,1}  ->  Y[row][member], then back to lane = member
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
          for (int nt = 0; nt < 4; nt++) {
            double2 v; v.x = c[mt][nt][0]; v.y = c[mt][nt][1];
            *reinterpret_cast<double2 *>(&Y[(mt * 8 + lane / 4) * 32 + nt * 8 + 2 * (lane % 4)]) = v;
          }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < 9; r++) { const double y = Y[r * 32 + lane]; acc += y; u[r] = fma(y, 1e-9, u[r]); }
        __syncwarp();
      }
    }
  }
  if (acc == 12345.6789) sink[0] = acc;
}

template <int MODE> double run(int iters, const double *uin, double *sink) {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int blocks = sms * 4, threads = 128;            // 16 warps per SM, like the marching kernels' 12-16
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<blocks, threads>>>(iters / 8 + 1, uin, sink);
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(iters, uin, sink);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return (double)iters * blocks * (threads / 32) / (ms * 1e-3);   // 32-stencil sets per second
}

int main() {
  double hB[9][9];
  for (int r = 0; r < 9; r++) for (int s = 0; s < 9; s++) hB[r][s] = 0.01 * ((r * 7 + s * 3) % 11) - 0.05;
  cudaMemcpyToSymbol(cB, hB, sizeof(hB));
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const size_t n = (size_t)sms * 4 * 128 * 9;
  double *uin, *sink; cudaMalloc(&uin, n * sizeof(double)); cudaMalloc(&sink, 8);
  double *h = new double[n];
  for (size_t i = 0; i < n; i++) h[i] = 1.0 + 1e-3 * (i % 97);
  cudaMemcpy(uin, h, n * sizeof(double), cudaMemcpyHostToDevice);
  const int iters = 20000;
  const double r0 = run<0>(iters, uin, sink), r1 = run<1>(iters, uin, sink), r2 = run<2>(iters, uin, sink), r3 = run<3>(iters, uin, sink);
  printf("ord-9 bridge mat-vec, 32 stencils per set, %d SMs, 16 warps/SM\n", sms);
  printf("mode 0 scalar full rows (81 DFMA/lane)        : %8.3f G sets/s   time vs mode 1: %.2fx   (%.2f T FMA/s)\n", r0 / 1e9, r1 / r0, r0 * 32 * 81 / 1e12);
  printf("mode 1 scalar even/odd rows (41 FMA + 8 add)   : %8.3f G sets/s   time vs mode 1: %.2fx\n", r1 / 1e9, 1.0);
  printf("mode 2 DMMA m8n8k4, operands in place (24/warp): %8.3f G sets/s   time vs mode 1: %.2fx   (%.2f T FMA/s incl. padding)\n", r2 / 1e9, r1 / r2, r2 * 24 * 256 / 1e12);
  printf("mode 3 DMMA + layout exchange through smem     : %8.3f G sets/s   time vs mode 1: %.2fx\n", r3 / 1e9, r1 / r3);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
