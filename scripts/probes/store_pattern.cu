// Store-pattern ceiling of the 4-state traversal kernel (profiles/README.md, "ceiling check"): the kernel writes
// every parent CLV once, a warp at a time in 4 KB pieces (4 x STG.256 per lane, 8 full 128-byte lines per
// instruction), walking NOPS different 128 MB buffers per piece.  This probe issues exactly those stores and nothing
// else, with the kernel's launch shape (2 CTAs of 8 warps per SM), and prints the achieved GB/s next to a plain
// linear fill of the same bytes.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/store_pattern store_pattern.cu && /tmp/store_pattern
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void stg256(double * p, double a, double b, double c, double d)
{
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}

// mode 0: op-major per warp (the traversal kernel's order); mode 1: linear fill with the same instruction
__global__ void __launch_bounds__(256, 2) pattern(double * base, unsigned long long buf_doubles, int nops,
                                                  unsigned long long sites, int mode)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned long long gwarp = (unsigned long long)blockIdx.x * 8 + warp, nwarps = (unsigned long long)gridDim.x * 8;
  const unsigned long long nchunks = sites / 32; /* 32 sites x 128 B = 4 KB per warp and buffer */
  const double v = (double)lane;
  if (mode == 0)
  {
    for (unsigned long long chunk = gwarp; chunk < nchunks; chunk += nwarps)
    {
      const unsigned long long s0 = chunk * 32 + (unsigned long long)(lane / 4) * 4;
      double * p = base + (s0 * 4 + (lane & 3)) * 4;
      for (int o = 0; o < nops; ++o)
      {
        double * q = p + (unsigned long long)o * buf_doubles;
#pragma unroll
        for (int r = 0; r < 4; ++r) stg256(q + 16 * r, v, v + o, v + r, v);
      }
    }
  }
  else
  {
    const unsigned long long total = (unsigned long long)nops * buf_doubles / 4; /* 32-byte elements */
    for (unsigned long long e = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (unsigned long long)gridDim.x * blockDim.x)
      stg256(base + e * 4, v, v, v, v);
  }
}

int main()
{
  cudaDeviceProp pr;
  cudaGetDeviceProperties(&pr, 0);
  const unsigned long long sites = 1000000 / 32 * 32; /* C2: 1 M sites, 128 B per site and buffer */
  const unsigned long long buf_doubles = sites * 16;
  const int nops = 256; /* 32.8 GB */
  double * d;
  if (cudaMalloc(&d, (size_t)nops * buf_doubles * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode)
    for (int rep = 0; rep < 3; ++rep)
    {
      cudaEventRecord(e0);
      pattern<<<pr.multiProcessorCount * 2, 256>>>(d, buf_doubles, nops, sites, mode);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      if (rep)
        printf("{\"pattern\": \"%s\", \"GB\": %.1f, \"ms\": %.3f, \"GBps\": %.1f}\n",
               mode == 0 ? "traversal order (per warp: 4 KB into each of 256 buffers)" : "linear fill, same instruction",
               nops * buf_doubles * 8 / 1e9, ms, nops * buf_doubles * 8 / 1e6 / ms);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
