// FP64 tensor-pipe probe for the 20-state kernel's roofline ("measure a DGEMM/DMMA peak on the
// box"): register-only mma.sync.m8n8k4.f64 chains, no memory traffic.  Prints the achieved TFLOP/s for a sweep
// of resident warps per SM and independent accumulator chains per warp, and the latency of a dependent chain.
// build + run on the GPU box:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/dmma_peak dmma_peak.cu && /tmp/dmma_peak
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double & c0, double & c1, double a, double b)
{
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

template <int CHAINS>
__global__ void probe(double * out, int iters, long long * cycles)
{
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  double c[CHAINS][2];
#pragma unroll
  for (int j = 0; j < CHAINS; ++j) c[j][0] = c[j][1] = 0.0;
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i)
#pragma unroll
    for (int j = 0; j < CHAINS; ++j) dmma(c[j][0], c[j][1], a, b);
  const long long t1 = clock64();
  double s = 0.0;
#pragma unroll
  for (int j = 0; j < CHAINS; ++j) s += c[j][0] + c[j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0;
}

template <int CHAINS>
static void run(int warps_per_sm, int sms, double * d_out, long long * d_cyc)
{
  const int iters = 20000;
  const int threads = warps_per_sm >= 8 ? 256 : warps_per_sm * 32;
  const int blocks = sms * (warps_per_sm * 32 / threads);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  probe<CHAINS><<<blocks, threads>>>(d_out, 100, d_cyc);
  cudaEventRecord(e0);
  probe<CHAINS><<<blocks, threads>>>(d_out, iters, d_cyc);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  long long cyc = 0;
  cudaMemcpy(&cyc, d_cyc, sizeof(cyc), cudaMemcpyDeviceToHost);
  const double n = (double)blocks * (threads / 32) * iters * CHAINS;
  printf("{\"warps_per_sm\": %d, \"chains_per_warp\": %d, \"TFLOPs_fp64\": %.2f, \"cycles_per_dmma_per_warp\": %.1f}\n",
         warps_per_sm, CHAINS, n * 512.0 / (ms * 1e-3) / 1e12, (double)cyc / ((double)iters * CHAINS));
}

int main()
{
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  double * d_out;
  long long * d_cyc;
  cudaMalloc(&d_out, sizeof(double) * p.multiProcessorCount * 2048);
  cudaMalloc(&d_cyc, sizeof(long long));
  printf("{\"device\": \"%s\", \"sms\": %d}\n", p.name, p.multiProcessorCount);
  const int ws[] = {4, 8, 16, 32};
  run<1>(4, p.multiProcessorCount, d_out, d_cyc); /* one warp per scheduler, one chain: the dependent-issue latency */
  for (int w : ws)
  {
    run<2>(w, p.multiProcessorCount, d_out, d_cyc);
    run<4>(w, p.multiProcessorCount, d_out, d_cyc);
    run<8>(w, p.multiProcessorCount, d_out, d_cyc);
  }
  return 0;
}
