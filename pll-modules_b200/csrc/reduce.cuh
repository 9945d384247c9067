/*
 * reduce.cuh — deterministic grid reduction used by the logL and derivative kernels.
 *
 * warp shuffle -> block (shared memory) -> one partial per CTA in HBM -> the CTA that takes the
 * last ticket sums the partials in a fixed tree order and writes the result both to the device result
 * slot (for a following NCCL all-reduce) and to mapped pinned host memory (so the host needs no
 * extra copy after the stream sync).  For a fixed grid size the summation order is fixed, so
 * repeated evaluations are bit-identical (the reference sums sequentially over sites,
 * core_likelihood.c:1484-1488; agreement is to rounding, ~1e-15 relative).
 */
#ifndef PLLB200_REDUCE_CUH_
#define PLLB200_REDUCE_CUH_

#include <cuda_runtime.h>

struct ReduceOut
{
  double * partials;       /* [N][gridDim.x] */
  unsigned int * ticket;
  double * result_dev;     /* slot base: N doubles */
  double * result_host;    /* mapped host alias of the same slot */
  /* completion tag: written to mapped host memory after the result (system-scope fence in between), so
   * the host can poll for a result instead of paying a stream synchronisation; null = a later kernel on
   * the stream finishes this slot (ascertainment-bias correction) */
  unsigned long long * tag_host;
  unsigned long long tag;
};

__device__ __forceinline__ void publish_tag(const ReduceOut & r)
{
  __threadfence_system();
  if (r.tag_host) *reinterpret_cast<volatile unsigned long long *>(r.tag_host) = r.tag;
}

template <int N>
__device__ __forceinline__ void grid_reduce_finish(double (&v)[N], const ReduceOut & r)
{
  __shared__ double s_red[N][32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int n = 0; n < N; ++n)
  {
    double x = v[n];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
    if (lane == 0) s_red[n][warp] = x;
  }
  __syncthreads();
  if (warp == 0)
  {
#pragma unroll
    for (int n = 0; n < N; ++n)
    {
      double x = lane < nwarps ? s_red[n][lane] : 0.0;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
      if (lane == 0) r.partials[(size_t)n * gridDim.x + blockIdx.x] = x;
    }
    if (lane == 0)
    {
      __threadfence();
      const unsigned int t = atomicAdd(r.ticket, 1u);
      s_last = (t == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  /* the last CTA: every thread takes partials b = tid, tid + blockDim, ... (one L2 round trip for up to blockDim
   * CTAs; a single warp walking gridDim / 32 dependent loads cost ~3.6 us at 296 CTAs), then the block tree above */
#pragma unroll
  for (int n = 0; n < N; ++n)
  {
    double x = 0.0;
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x)
      x += __ldcg(r.partials + (size_t)n * gridDim.x + b);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
    if (lane == 0) s_red[n][warp] = x;
  }
  __syncthreads();
  if (warp == 0)
  {
#pragma unroll
    for (int n = 0; n < N; ++n)
    {
      double x = lane < nwarps ? s_red[n][lane] : 0.0;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
      if (lane == 0)
      {
        r.result_dev[n] = x;
        r.result_host[n] = x;
      }
    }
    if (lane == 0)
    {
      *r.ticket = 0;
      publish_tag(r);
    }
  }
}

#endif
