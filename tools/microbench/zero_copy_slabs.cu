// Microbenchmark for the host-buffer path: member slabs of (rows, nens) fields moved between pinned host
// memory and the device (a) by the copy engines (cudaMemcpy2DAsync, what awfl_time_step_host does) and (b) by a small
// kernel that loads / stores the mapped host memory directly, with the opposite direction busy on a copy engine at the
// same time. Prints GB/s per direction.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/zc tools/microbench/zero_copy_slabs.cu && /tmp/zc
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

// slab [e0, e0+ne) of nf fields of rows x nens doubles, 16 bytes per thread and iteration, grid-stride
__global__ void k_slab_copy(double2 *__restrict__ dst, const double2 *__restrict__ src, int nf, long rows, int nens2, int e02, int ne2) {
  const long per_field = rows * ne2, total = per_field * nf;
  for (long t = (long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long)gridDim.x * blockDim.x) {
    const long f = t / per_field, r = (t - f * per_field) / ne2;
    const int e = (int)(t - f * per_field - r * ne2);
    const long idx = (f * rows + r) * nens2 + e02 + e;
    dst[idx] = src[idx];
  }
}

int main(int argc, char **argv) {
  const int nf = 6, nens = 4096, P = argc > 1 ? atoi(argv[1]) : 13;
  const long rows = 58 * 64;
  const size_t bytes = sizeof(double) * nf * rows * nens;
  double *h_in, *h_out, *d_in, *d_out;
  CK(cudaHostAlloc((void **)&h_in, bytes, cudaHostAllocMapped));
  // argv[2] = 1: in place, like awfl_time_step_host (results overwrite the inputs: the two directions touch the SAME host
  // rows, other columns, at the same time -- a slab's results go out while a later slab's inputs come in)
  const bool inplace = argc > 2 && atoi(argv[2]) == 1;
  if (inplace) h_out = h_in; else
  CK(cudaHostAlloc((void **)&h_out, bytes, cudaHostAllocMapped));
  CK(cudaMalloc((void **)&d_in, bytes));
  CK(cudaMalloc((void **)&d_out, bytes));
  for (size_t i = 0; i < bytes / 8; i++) h_in[i] = (double)i;
  CK(cudaMemset(d_out, 1, bytes));
  cudaStream_t s_in, s_out;
  CK(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
  cudaEvent_t a0, a1, b0, b1;
  CK(cudaEventCreate(&a0)); CK(cudaEventCreate(&a1)); CK(cudaEventCreate(&b0)); CK(cudaEventCreate(&b1));
  const int tiles = nens / 32;
  const size_t pitch = sizeof(double) * nens;
  // mode bits: 1 = H2D by kernel, 2 = D2H by kernel; dirs: 1 = H2D only, 2 = D2H only, 3 = both at once
  for (int grid : {0, 32, 148, 592}) {
    for (int dirs = 1; dirs <= 3; dirs++) {
      for (int rep = 0; rep < 3; rep++) {
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(a0, s_in)); CK(cudaEventRecord(b0, s_out));
        for (int s = 0; s < P; s++) {
          const int t0 = (int)((long)tiles * s / P), t1 = (int)((long)tiles * (s + 1) / P);
          const int e0 = 32 * t0, ne = 32 * (t1 - t0);
          if (dirs & 1) {
            if (grid == 0) for (int v = 0; v < nf; v++)
              CK(cudaMemcpy2DAsync(d_in + v * rows * nens + e0, pitch, h_in + v * rows * nens + e0, pitch, sizeof(double) * ne, rows, cudaMemcpyHostToDevice, s_in));
            else k_slab_copy<<<grid, 256, 0, s_in>>>((double2 *)d_in, (const double2 *)h_in, nf, rows, nens / 2, e0 / 2, ne / 2);
          }
          if ((dirs & 2) && inplace && grid == 0 && dirs == 3) {
            // the results of slab s-3 (the pipeline's depth) leave while slab s arrives
            const int so = s - 3;
            if (so >= 0) {
              const int u0 = (int)((long)tiles * so / P), u1 = (int)((long)tiles * (so + 1) / P);
              for (int v = 0; v < nf; v++)
                CK(cudaMemcpy2DAsync(h_out + v * rows * nens + 32 * u0, pitch, d_out + v * rows * nens + 32 * u0, pitch, sizeof(double) * 32 * (u1 - u0), rows, cudaMemcpyDeviceToHost, s_out));
            }
          } else if (dirs & 2) {
            if (grid == 0) for (int v = 0; v < nf; v++)
              CK(cudaMemcpy2DAsync(h_out + v * rows * nens + e0, pitch, d_out + v * rows * nens + e0, pitch, sizeof(double) * ne, rows, cudaMemcpyDeviceToHost, s_out));
            else k_slab_copy<<<grid, 256, 0, s_out>>>((double2 *)h_out, (const double2 *)d_out, nf, rows, nens / 2, e0 / 2, ne / 2);
          }
        }
        CK(cudaEventRecord(a1, s_in)); CK(cudaEventRecord(b1, s_out));
        CK(cudaDeviceSynchronize());
        CK(cudaGetLastError());
        float ma = 0.f, mb = 0.f;
        CK(cudaEventElapsedTime(&ma, a0, a1)); CK(cudaEventElapsedTime(&mb, b0, b1));
        if (rep == 2)
          printf("%-22s P=%d dirs=%d  H2D %6.2f ms %5.1f GB/s   D2H %6.2f ms %5.1f GB/s\n", grid == 0 ? "copy engines (2-D)" : "kernel, mapped host", P, dirs,
                 ma, (dirs & 1) ? bytes / ma * 1e-6 : 0.0, mb, (dirs & 2) ? bytes / mb * 1e-6 : 0.0);
      }
    }
    if (grid) printf("   (grid %d x 256 threads)\n", grid);
  }
  // mixed: H2D by the copy engine, D2H by the kernel (the candidate for awfl_time_step_host), and the reverse
  for (int grid : {32, 148}) {
    for (int which = 0; which < 2; which++) {
      for (int rep = 0; rep < 3; rep++) {
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(a0, s_in)); CK(cudaEventRecord(b0, s_out));
        for (int s = 0; s < P; s++) {
          const int t0 = (int)((long)tiles * s / P), t1 = (int)((long)tiles * (s + 1) / P);
          const int e0 = 32 * t0, ne = 32 * (t1 - t0);
          if (which == 0) {
            for (int v = 0; v < nf; v++)
              CK(cudaMemcpy2DAsync(d_in + v * rows * nens + e0, pitch, h_in + v * rows * nens + e0, pitch, sizeof(double) * ne, rows, cudaMemcpyHostToDevice, s_in));
            k_slab_copy<<<grid, 256, 0, s_out>>>((double2 *)h_out, (const double2 *)d_out, nf, rows, nens / 2, e0 / 2, ne / 2);
          } else {
            k_slab_copy<<<grid, 256, 0, s_in>>>((double2 *)d_in, (const double2 *)h_in, nf, rows, nens / 2, e0 / 2, ne / 2);
            for (int v = 0; v < nf; v++)
              CK(cudaMemcpy2DAsync(h_out + v * rows * nens + e0, pitch, d_out + v * rows * nens + e0, pitch, sizeof(double) * ne, rows, cudaMemcpyDeviceToHost, s_out));
          }
        }
        CK(cudaEventRecord(a1, s_in)); CK(cudaEventRecord(b1, s_out));
        CK(cudaDeviceSynchronize());
        float ma = 0.f, mb = 0.f;
        CK(cudaEventElapsedTime(&ma, a0, a1)); CK(cudaEventElapsedTime(&mb, b0, b1));
        if (rep == 2)
          printf("%-34s grid %3d  H2D %6.2f ms %5.1f GB/s   D2H %6.2f ms %5.1f GB/s\n", which == 0 ? "H2D copy engine + D2H kernel" : "H2D kernel + D2H copy engine",
                 grid, ma, bytes / ma * 1e-6, mb, bytes / mb * 1e-6);
      }
    }
  }
  // check that the kernel-written host buffer holds the device pattern
  const unsigned char *hb = (const unsigned char *)h_out;
  size_t bad = 0;
  if (!inplace) for (size_t i = 0; i < bytes; i += 4097) bad += hb[i] != 1;
  printf("host buffer check: %zu mismatches\n", bad);
  return bad != 0;
}
