// tools/ubench.cu -- single-warp / single-CTA latency microbenchmarks on B200 for the FP64 chain ops the
// batch kernel's critical path is made of.  nvcc -O3 -arch=sm_100a tools/ubench.cu -o /tmp/ubench
#include <cstdio>
#include <cuda_runtime.h>
#define N 4096
__global__ void k_dfma_dep(double *out, double a, double b) { double x = out[threadIdx.x]; long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) x = fma(x, a, b);
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("dependent DFMA: %.1f cycles each (%d warps)\n", double(t1 - t0) / N, blockDim.x / 32); }
__global__ void k_dfma_ind(double *out, double a, double b) { double x[8]; for (int j = 0; j < 8; j++) x[j] = out[threadIdx.x + j]; long long t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) x[j] = fma(x[j], a, b); }
  long long t1 = clock64(); double s = 0; for (int j = 0; j < 8; j++) s += x[j]; out[threadIdx.x] = s;
  if (threadIdx.x == 0) printf("8 independent DFMA chains: %.1f cycles per DFMA (%d warps)\n", double(t1 - t0) / N / 8, blockDim.x / 32); }
__global__ void k_ddiv_dep(double *out, double a) { double x = out[threadIdx.x]; long long t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; i++) x = a / x + 1.0;
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("dependent DDIV+DADD: %.1f cycles each\n", double(t1 - t0) / N); }
__global__ void k_drcp_dep(double *out) { double x = out[threadIdx.x]; long long t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; i++) x = __drcp_rn(x) + 1.0;
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("dependent DRCP+DADD: %.1f cycles each\n", double(t1 - t0) / N); }
__global__ void k_lds_dep(double *out) { __shared__ int s[1024]; for (int i = threadIdx.x; i < 1024; i += blockDim.x) s[i] = (i * 33 + 7) & 1023; __syncthreads();
  int p = threadIdx.x; long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) p = s[p];
  long long t1 = clock64(); out[threadIdx.x] = p; if (threadIdx.x == 0) printf("dependent LDS: %.1f cycles each\n", double(t1 - t0) / N); }
__global__ void k_sts_lds(double *out) { __shared__ double s[1024]; double x = out[threadIdx.x]; s[threadIdx.x] = x; __syncthreads(); long long t0 = clock64();
#pragma unroll 8
  for (int i = 0; i < N; i++) { s[threadIdx.x] = x; __syncwarp(); x = s[threadIdx.x ^ 1] + 1.0; __syncwarp(); }
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("STS -> LDS (other lane) -> DADD round trip: %.1f cycles\n", double(t1 - t0) / N); }
__global__ void k_bar(double *out) { long long t0 = clock64();
#pragma unroll 8
  for (int i = 0; i < N; i++) __syncthreads();
  long long t1 = clock64(); if (threadIdx.x == 0) { printf("__syncthreads (%d warps): %.1f cycles each\n", blockDim.x / 32, double(t1 - t0) / N); out[0] = 1; } }
__global__ void k_bar_exchange(double *out) { __shared__ double s[128]; double x = out[threadIdx.x]; long long t0 = clock64();
  for (int i = 0; i < N; i++) { s[threadIdx.x] = x; __syncthreads(); x = fma(s[(threadIdx.x + 32) & 127], 1.0000001, 0.5); __syncthreads(); }
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("STS, bar, LDS, DFMA, bar (%d warps): %.1f cycles per round\n", blockDim.x / 32, double(t1 - t0) / N); }
__global__ void k_shfl(double *out) { double x = out[threadIdx.x]; long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31) + 1.0;
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("dependent SHFL(double)+DADD: %.1f cycles each\n", double(t1 - t0) / N); }
__global__ void k_iadd_dep(double *out) { int x = (int)out[threadIdx.x]; long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) x = x * 3 + i;
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("dependent IMAD: %.1f cycles each\n", double(t1 - t0) / N); }
int main() { double *d; cudaMalloc(&d, 8192 * 8); cudaMemset(d, 0, 8192 * 8); double h[1024]; for (int i = 0; i < 1024; i++) h[i] = 1.5 + i * 1e-3; cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice);
  k_dfma_dep<<<1, 32>>>(d, 0.999999, 1e-7); k_dfma_dep<<<1, 128>>>(d, 0.999999, 1e-7); k_dfma_ind<<<1, 32>>>(d, 0.999999, 1e-7); k_dfma_ind<<<1, 128>>>(d, 0.999999, 1e-7); k_dfma_ind<<<1, 512>>>(d, 0.999999, 1e-7);
  cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice); k_ddiv_dep<<<1, 32>>>(d, 1.7); cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice); k_drcp_dep<<<1, 32>>>(d);
  k_lds_dep<<<1, 32>>>(d); cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice); k_sts_lds<<<1, 32>>>(d); k_bar<<<1, 128>>>(d); k_bar<<<1, 256>>>(d);
  cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice); k_bar_exchange<<<1, 128>>>(d); k_shfl<<<1, 32>>>(d); k_iadd_dep<<<1, 32>>>(d);
  cudaDeviceSynchronize(); printf("%s\n", cudaGetErrorString(cudaGetLastError())); return 0; }
