// tools/ubench2.cu -- where do the cycles of the backward solve's 4-row step go?  Single warp, shared memory.
// nvcc -O3 -arch=sm_100a tools/ubench2.cu -o /tmp/ubench2
#include <cstdio>
#include <cuda_runtime.h>
#define REP 2000
template <int OFF> __device__ __forceinline__ double lds(unsigned a) { double v; asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(a), "n"(OFF)); return v; }
template <int MODE> __global__ void k_block(double *out, int nwarm)
{
    __shared__ double L[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) L[i] = 1e-3 * ((i * 37) % 101 - 50);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned Ls = (unsigned)__cvta_generic_to_shared(L);
    double v = 1.0 + lane;
    long long t0 = clock64();
    for (int rep = 0; rep < REP; rep++) {
        int i0 = 19, g = 9;
        unsigned am = Ls + 8u * (unsigned)(lane * 20 + i0 + 64), a1 = Ls + 8u * 900u;
        const int lim = lane;
#pragma unroll 1
        for (; i0 >= 3; i0 -= 4) {
            const unsigned a2 = a1 - 8u * (unsigned)(g + 1), a3 = a2 - 8u * (unsigned)(g + 2);
            const double l01 = lds<0>(a1), l02 = lds<0>(a2), l12 = lds<-8>(a2);
            const double l03 = lds<0>(a3), l13 = lds<-8>(a3), l23 = lds<-16>(a3);
            double m0 = lds<0>(am), m1 = lds<-8>(am), m2 = lds<-16>(am), m3 = lds<-24>(am);
            m0 = lim < i0 ? m0 : 0.0; m1 = lim < i0 - 1 ? m1 : 0.0; m2 = lim < i0 - 2 ? m2 : 0.0; m3 = lim < i0 - 3 ? m3 : 0.0;
            double x0, v1, v2, v3;
            if (MODE == 1) { x0 = v; v1 = v + 1.0; v2 = v + 2.0; v3 = v + 3.0; }          // no shuffles
            else { x0 = __shfl_sync(0xffffffffu, v, i0); v1 = __shfl_sync(0xffffffffu, v, i0 - 1); v2 = __shfl_sync(0xffffffffu, v, i0 - 2); v3 = __shfl_sync(0xffffffffu, v, i0 - 3); }
            const double x1 = fma(-l01, x0, v1);
            const double x2 = fma(-l12, x1, fma(-l02, x0, v2));
            const double x3 = fma(-l23, x2, fma(-l13, x1, fma(-l03, x0, v3)));
            if (MODE == 2) { v = fma(-m0, x0, v) + fma(-m1, x1, 0.0) + fma(-m2, x2, 0.0) + fma(-m3, x3, 0.0); }
            else { v = fma(-m0, x0, v); v = fma(-m1, x1, v); v = fma(-m2, x2, v); v = fma(-m3, x3, v); }
            a1 = a3 - 8u * (unsigned)(2 * g + 7) - 32u; am -= 32u; g += 4;
        }
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) printf("mode %d: %.1f cycles per 4-row step (%d warps)\n", MODE, double(t1 - t0) / REP / 5, blockDim.x / 32);
}
// software-pipelined: the L entries of step k+1 are loaded during the chain of step k
template <int MODE> __global__ void k_block_pf(double *out)
{
    __shared__ double L[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) L[i] = 1e-3 * ((i * 37) % 101 - 50);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned Ls = (unsigned)__cvta_generic_to_shared(L);
    double v = 1.0 + lane;
    long long t0 = clock64();
    for (int rep = 0; rep < REP; rep++) {
        int i0 = 19, g = 9;
        unsigned am = Ls + 8u * (unsigned)(lane * 20 + i0 + 64), a1 = Ls + 8u * 900u;
        const int lim = lane;
        unsigned a2 = a1 - 8u * (unsigned)(g + 1), a3 = a2 - 8u * (unsigned)(g + 2);
        double l01 = lds<0>(a1), l02 = lds<0>(a2), l12 = lds<-8>(a2), l03 = lds<0>(a3), l13 = lds<-8>(a3), l23 = lds<-16>(a3);
        double m0 = lds<0>(am), m1 = lds<-8>(am), m2 = lds<-16>(am), m3 = lds<-24>(am);
#pragma unroll 1
        for (; i0 >= 3; i0 -= 4) {
            // next step's addresses and loads (one step past the end reads unrelated entries that are never used)
            a1 = a3 - 8u * (unsigned)(2 * g + 7) - 32u; am -= 32u; g += 4;
            a2 = a1 - 8u * (unsigned)(g + 1); a3 = a2 - 8u * (unsigned)(g + 2);
            const double n01 = lds<0>(a1), n02 = lds<0>(a2), n12 = lds<-8>(a2), n03 = lds<0>(a3), n13 = lds<-8>(a3), n23 = lds<-16>(a3);
            const double q0 = lds<0>(am), q1 = lds<-8>(am), q2 = lds<-16>(am), q3 = lds<-24>(am);
            m0 = lim < i0 ? m0 : 0.0; m1 = lim < i0 - 1 ? m1 : 0.0; m2 = lim < i0 - 2 ? m2 : 0.0; m3 = lim < i0 - 3 ? m3 : 0.0;
            double x0, v1, v2, v3;
            if (MODE == 1) { x0 = v; v1 = v + 1.0; v2 = v + 2.0; v3 = v + 3.0; }
            else { x0 = __shfl_sync(0xffffffffu, v, i0); v1 = __shfl_sync(0xffffffffu, v, i0 - 1); v2 = __shfl_sync(0xffffffffu, v, i0 - 2); v3 = __shfl_sync(0xffffffffu, v, i0 - 3); }
            const double x1 = fma(-l01, x0, v1);
            const double x2 = fma(-l12, x1, fma(-l02, x0, v2));
            const double x3 = fma(-l23, x2, fma(-l13, x1, fma(-l03, x0, v3)));
            v = fma(-m0, x0, v); v = fma(-m1, x1, v); v = fma(-m2, x2, v); v = fma(-m3, x3, v);
            l01 = n01; l02 = n02; l12 = n12; l03 = n03; l13 = n13; l23 = n23; m0 = q0; m1 = q1; m2 = q2; m3 = q3;
        }
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) printf("prefetched, mode %d: %.1f cycles per 4-row step (%d warps)\n", MODE, double(t1 - t0) / REP / 5, blockDim.x / 32);
}
__global__ void k_shfl_lat(double *out) { double x = out[threadIdx.x]; int idx = (threadIdx.x * 7 + 3) & 31; long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 4096; i++) { x = __shfl_sync(0xffffffffu, x, idx) + 1.0; idx = (idx + 5) & 31; }
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("dependent SHFL(double, register index)+DADD, rolled: %.1f cycles each\n", double(t1 - t0) / 4096); }
__global__ void k_dfma_chain(double *out) { double x = out[threadIdx.x], a = out[threadIdx.x + 32]; long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 4096; i++) { x = fma(-a, x, 1.0); x = fma(-a, x, 1.0); x = fma(-a, x, 1.0); x = fma(-a, x, 1.0); }
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("4 dependent DFMA per rolled iteration: %.1f cycles per DFMA\n", double(t1 - t0) / 4096 / 4); }
__global__ void k_lds_fma(double *out) { __shared__ double s[1024]; for (int i = threadIdx.x; i < 1024; i += blockDim.x) s[i] = 1e-3 * i; __syncthreads();
  const unsigned b = (unsigned)__cvta_generic_to_shared(s) + 8u * threadIdx.x; double x = out[threadIdx.x]; long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 4096; i++) { const double a0 = lds<0>(b), a1 = lds<256>(b), a2 = lds<512>(b), a3 = lds<768>(b); x = fma(-a0, x, 1.0); x = fma(-a1, x, 1.0); x = fma(-a2, x, 1.0); x = fma(-a3, x, 1.0); }
  long long t1 = clock64(); out[threadIdx.x] = x; if (threadIdx.x == 0) printf("4 LDS + 4 dependent DFMA per rolled iteration: %.1f cycles per iteration\n", double(t1 - t0) / 4096); }
int main() { double *d; cudaMalloc(&d, 8192 * 8); double h[1024]; for (int i = 0; i < 1024; i++) h[i] = 0.5 + i * 1e-3; cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice);
  k_block<0><<<1, 32>>>(d, 0); k_block<1><<<1, 32>>>(d, 0); k_block<2><<<1, 32>>>(d, 0); k_block<0><<<1, 128>>>(d, 0); k_block_pf<0><<<1, 32>>>(d); k_block_pf<1><<<1, 32>>>(d); k_block_pf<0><<<1, 128>>>(d);
  cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice); k_shfl_lat<<<1, 32>>>(d); k_dfma_chain<<<1, 32>>>(d); k_lds_fma<<<1, 32>>>(d);
  cudaDeviceSynchronize(); printf("%s\n", cudaGetErrorString(cudaGetLastError())); return 0; }
