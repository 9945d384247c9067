/*
 * k_newton.cu — the whole Newton-Raphson iteration on ONE branch length in ONE persistent kernel.
 *
 * The reference's inner loop (pllmod_opt_minimize_newton_multi, src/optimize/opt_algorithms.c:133-262,
 * driven from pll_optimize.c:1395-1560) calls pll_compute_likelihood_derivatives once per iteration:
 * on a GPU that is one launch + one host wait per iteration (3-30 per branch), and branch-length
 * optimisation is what an SPR round spends 85 % of its time in (profiles/README.md).  Here a
 * cooperative grid keeps the bracket [xl, xh], the iterate and the convergence tests on the device:
 *
 *   repeat:  every CTA rebuilds the diagonal table diag[k][j] = (e, l e, l^2 e), e = exp(l t),
 *            l = eigenvalue_j * rate_k / (1 - pinv_k)               (core_derivatives.c:757-772)
 *            grid-stride pass over the sumtable -> per-CTA partial (d_f, dd_f)     (A.7)
 *            grid barrier; every CTA sums the partials in index order (deterministic, identical
 *            in all CTAs) and applies the SAME damped Newton update to its private copy of the state
 *   until converged / iteration limit / non-finite derivative.
 *
 * One launch and one host wait per branch; the sumtable (one CLV-sized buffer) is re-read from L2
 * when it fits.  Conditions for this path (checked by the host, otherwise the per-iteration path
 * runs): one local partition, no cross-rank reduction callback, no ascertainment bias, rate
 * categories a power of two <= 32, K * states <= 512.
 */
#include "internal.h"
#include "devutil.cuh"
#include "aa4_pass.cuh"
#include "reduce.cuh"

#include <algorithm>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

#define NW_THREADS 256
#define NW_MAXK 32
#define NW_MAXRANKS 8
#define NW_MAX_DIAG 512 /* K * states */
#define NW_DMMA_MAXE 256 /* MODE 5: K * states_padded */

/* status codes returned in result[1] */
#define NW_OK 0.0
#define NW_LIMIT 1.0
#define NW_BAD_DERIV 2.0

struct NewtonArgs
{
  const double * sum;
  const unsigned int * weights;
  const int * invariant; /* may be null */
  const double * eigen;
  unsigned long long eigen_stride;
  unsigned int sites, states, sp, K, kshift, any_pinv;
  unsigned int model_idx[NW_MAXK];
  double w[NW_MAXK], pinv[NW_MAXK], rate[NW_MAXK];
  double xmin, xmax, x0, tol, scale; /* scale: brlen scaler s (df *= s, ddf *= s^2, t = s x) */
  unsigned int max_iters;
  double * partials; /* [2 buffers][2][gridDim.x] */
  double * result_dev;
  double * result_host;
  unsigned long long * tag_host; /* completion tag in mapped host memory, written after the result */
  unsigned long long tag;
  /* queued mode (a whole smoothing pass without a host wait): {x, x * scale, status, iterations} of this
   * branch in device memory; the start value comes from there when the branch was already visited */
  double * brlen_io;
  int x0_from_device, revert_on_limit;
  /* newton_dna4_tma_kernel only: 1 / 2 = one pass at x0 (backwards / forwards over the table), result {d_f, dd_f}:
   * pll_compute_likelihood_derivatives on the TMA ring (k_derivatives.cu) */
  int eval_only;
  unsigned int * ticket; /* eval_only: last-CTA reduction (reduce.cuh) instead of a grid barrier */
  /* multi-GPU (site-sharded ranks, one process per GPU): every pass ends with an all-gather of the local
   * (d_f, dd_f) over NVLink peer memory instead of a host callback.  inbox[r] is rank r's exchange
   * buffer (own: local pointer, others: CUDA IPC mappings): [2 parities][8 ranks]{f, df, seq, -} then the
   * local mailbox [2]{f, df, seq, -} and the running pass counter. */
  int rank, nranks;
  double * inbox[NW_MAXRANKS];
};

#define NW_XCHG_MAIL 64  /* doubles: offset of the local mailbox inside an exchange buffer */
#define NW_XCHG_COUNTER 72
#define NW_XCHG_DOUBLES 80
#define NW_SPIN_LIMIT 40000000ull /* tens of seconds: a rank that never shows up must not hang the GPU */

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long * p)
{
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long * p, unsigned long long v)
{
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long * p)
{
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long * p, unsigned long long v)
{
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

/* Called by ONE thread per CTA with this rank's totals: CTA 0 pushes them into every rank's inbox, waits
 * for all ranks' values in its own inbox, adds them in rank order (identical on every GPU) and leaves the
 * result in the local mailbox, where the other CTAs pick it up.  `pass` is the running pass number (same
 * on all ranks, alternating parity => a slot is rewritten only after everybody has read it).  Returns
 * false if a peer did not answer. */
template <class Args> /* NewtonArgs or NewtonMultiArgs: needs rank, nranks, inbox[] */
__device__ __forceinline__ bool newton_allgather(const Args & a, unsigned long long pass, double & t1,
                                                 double & t2)
{
  const unsigned int par = (unsigned int)(pass & 1ull);
  double * own = a.inbox[a.rank];
  double * mail = own + NW_XCHG_MAIL + par * 4;
  unsigned long long spins = 0;
  if (blockIdx.x == 0)
  {
    for (int r = 0; r < a.nranks; ++r)
    {
      double * slot = a.inbox[r] + (par * NW_MAXRANKS + a.rank) * 4;
      slot[0] = t1;
      slot[1] = t2;
    }
    __threadfence_system();
    for (int r = 0; r < a.nranks; ++r)
      st_release_sys(reinterpret_cast<unsigned long long *>(a.inbox[r] + (par * NW_MAXRANKS + a.rank) * 4 + 2), pass);
    double g1 = 0.0, g2 = 0.0;
    for (int r = 0; r < a.nranks; ++r)
    {
      const double * slot = own + (par * NW_MAXRANKS + r) * 4;
      while (ld_acquire_sys(reinterpret_cast<const unsigned long long *>(slot + 2)) != pass)
        if (++spins > NW_SPIN_LIMIT) return false;
      g1 += slot[0];
      g2 += slot[1];
    }
    mail[0] = g1;
    mail[1] = g2;
    __threadfence();
    st_release_gpu(reinterpret_cast<unsigned long long *>(mail + 2), pass);
    t1 = g1;
    t2 = g2;
    return true;
  }
  while (ld_acquire_gpu(reinterpret_cast<const unsigned long long *>(mail + 2)) != pass)
    if (++spins > NW_SPIN_LIMIT) return false;
  t1 = mail[0];
  t2 = mail[1];
  return true;
}

/* After the grid barrier every CTA forms the same ordered sum of all CTAs' partials.  The CTA's first 256 threads
 * take one or two partials each — one L2 round trip — instead of one warp walking G / 32 dependent loads (296
 * CTAs: ten serial L2 latencies, 3.6 us of every pass).  red: [2][NW_THREADS / 32] shared doubles (the warp
 * partials of the pass, already consumed).  Returns the totals in thread 0 only. */
__device__ __forceinline__ void newton_sum_partials(const double * buf, unsigned int G, double (*red)[NW_THREADS / 32],
                                                    double & t1, double & t2)
{
  const unsigned int tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  double a1 = 0.0, a2 = 0.0;
  if (tid < NW_THREADS)
    for (unsigned int q = tid; q < G; q += NW_THREADS)
    {
      a1 += __ldcg(buf + q);
      a2 += __ldcg(buf + G + q);
    }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
  {
    a1 += __shfl_down_sync(0xffffffffu, a1, off);
    a2 += __shfl_down_sync(0xffffffffu, a2, off);
  }
  if (tid < NW_THREADS && lane == 0)
  {
    red[0][warp] = a1;
    red[1][warp] = a2;
  }
  __syncthreads();
  t1 = t2 = 0.0;
  if (tid == 0)
    for (int wi = 0; wi < NW_THREADS / 32; ++wi)
    {
      t1 += red[0][wi];
      t2 += red[1][wi];
    }
}

/* what the host does with a result (pll_optimize.c:1524-1560), for the queued mode: an unconverged or
 * failed branch keeps its old length, and so does a change below 1e-10 */
__device__ __forceinline__ void newton_publish(const NewtonArgs & a, double x, double xorig, double status,
                                               unsigned int iter)
{
  if (!a.brlen_io) return;
  if ((status == NW_LIMIT && a.revert_on_limit) || status == NW_BAD_DERIV) x = xorig;
  if (fabs(x - xorig) < 1e-10) x = xorig;
  a.brlen_io[0] = x;
  a.brlen_io[1] = x * a.scale;
  a.brlen_io[2] = status;
  a.brlen_io[3] = (double)iter;
}


/* MODE 0: any state count;  1: 4 states, any K;  2: 4 states, K = 4 (reduce-scatter over the rates);
 * 4: 20 states, K = 4, unpadded (protein Gamma-4: the pass is a [sites x 80] x [80 x 3] product on the FP64 tensor cores) */
template <int MODE>
__global__ void __launch_bounds__(NW_THREADS, (MODE == 2 || MODE == 5) ? 2 : 3) newton_kernel(NewtonArgs a)
{
  constexpr bool DNA = MODE == 1 || MODE == 2;
  __shared__ double s_aux[MODE == 5 ? 80 : 24]; /* MODE 4 / 5: w_k (1 - pinv_k) | invariant-site term per state */
  __shared__ double s_bt[MODE == 4 ? AA4_BT : (MODE == 5 ? NW_DMMA_MAXE * 8 : 1)]; /* B fragments of the tensor-core pass */
  cg::grid_group grid = cg::this_grid();
  __shared__ double s_diag[NW_MAX_DIAG * 3];
  __shared__ double s_red[2][NW_THREADS / 32];
  __shared__ double s_tot[2];

  const unsigned int K = a.K, s = a.states, sp = a.sp;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned int k = (unsigned int)lane & (K - 1); /* K is a power of two <= 32 */
  const unsigned int kshift = a.kshift;
  const double wk = a.w[k], pinv = a.pinv[k];
  const bool any_pinv = a.any_pinv != 0;
  const double * fr = a.eigen + (unsigned long long)a.model_idx[k] * a.eigen_stride + 2 * s * sp + sp;
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long Eround = (E + 31) / 32 * 32;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  const unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  /* same for all lanes of a warp: Eround and the warp's first index are multiples of 32 */
  const unsigned int n_it = base < Eround ? (unsigned int)((Eround - base + stride - 1) / stride) : 0u;
  constexpr int U = DNA ? 4 : 1; /* independent 32-byte loads in flight per thread */

  /* optimiser state, replicated in every thread (all threads see the same f, df) */
  double xl = a.xmin, xh = a.xmax;
  const double xorig = (a.brlen_io && a.x0_from_device) ? __ldcg(a.brlen_io) : a.x0;
  /* passes done by earlier launches (device-resident: the host does not know how many a launch needs) */
  const unsigned long long pass0 =
      a.nranks > 1 ? __ldcg(reinterpret_cast<const unsigned long long *>(a.inbox[a.rank] + NW_XCHG_COUNTER)) : 0ull;
  double x = fmax(fmin(xorig, a.xmax), a.xmin);
  const double dxmax = a.xmax / a.max_iters;
  unsigned int iter = 0;
  double status = NW_OK;
  double f = 0.0, df = 0.0;

  for (;;)
  {
    if (iter++ > a.max_iters)
    {
      status = NW_LIMIT;
      break;
    }
    /* diagonal table for the current branch length */
    const double t = a.scale * x;
    __syncthreads();
    for (unsigned int i = threadIdx.x; i < K * s; i += blockDim.x)
    {
      const unsigned int kk = i / s, j = i % s;
      const double ev = a.eigen[(unsigned long long)a.model_idx[kk] * a.eigen_stride + 2 * s * sp + j];
      const double ki = a.rate[kk] / (1.0 - a.pinv[kk]);
      const double e0 = exp(ev * ki * t);
      s_diag[i * 3 + 0] = e0;
      s_diag[i * 3 + 1] = ev * ki * e0;
      s_diag[i * 3 + 2] = ev * ki * ev * ki * e0;
      if (MODE == 2)
      {
        /* the K = 4 pass reads a second copy with w_k (1 - pinv_k) folded in (last 48 entries) */
        const double wf = a.w[kk] * (1.0 - a.pinv[kk]);
        double * d = s_diag + NW_MAX_DIAG * 3 - 48;
        d[i * 3 + 0] = wf * e0;
        d[i * 3 + 1] = wf * ev * ki * e0;
        d[i * 3 + 2] = wf * ev * ki * ev * ki * e0;
      }
    }
    __syncthreads();

    double acc1 = 0.0, acc2 = 0.0;
    const double * dg = s_diag + (size_t)k * s * 3;
    /* odd passes walk the table backwards: what the previous pass (or the sumtable kernel, which
     * writes ascending) touched last is still in the 126 MB L2 */
    const bool fwd = (iter & 1u) == 0;
    if (MODE == 4)
    {
      /* 20 states x 4 rates: the pass runs on the FP64 tensor cores (aa4_pass.cuh) */
      if (threadIdx.x < 4) s_aux[threadIdx.x] = a.w[threadIdx.x] * (1.0 - a.pinv[threadIdx.x]);
      if (threadIdx.x < 20)
      {
        double t = 0.0;
        for (unsigned int kk = 0; kk < 4; ++kk)
          t += a.w[kk] * a.pinv[kk] *
               a.eigen[(unsigned long long)a.model_idx[kk] * a.eigen_stride + 2 * 20 * 20 + 20 + threadIdx.x];
        s_aux[4 + threadIdx.x] = t;
      }
      __syncthreads();
      aa4_build_b(s_diag, s_aux, s_bt);
      __syncthreads();
      aa4_pass(a.sum, a.weights, a.invariant, (any_pinv && a.invariant) ? s_aux + 4 : nullptr, a.sites, s_bt,
               (unsigned long long)blockIdx.x * (NW_THREADS / 32) + warp,
               (unsigned long long)gridDim.x * (NW_THREADS / 32), fwd, lane, acc1, acc2);
    }
    else if (MODE == 5)
    {
      /* 17..64 states, K * states_padded a multiple of 16 and <= NW_DMMA_MAXE: the generic tensor-core pass */
      if (threadIdx.x < K) s_aux[threadIdx.x] = a.w[threadIdx.x] * (1.0 - a.pinv[threadIdx.x]);
      if (threadIdx.x < s)
      {
        double t = 0.0;
        for (unsigned int kk = 0; kk < K; ++kk)
          t += a.w[kk] * a.pinv[kk] * a.eigen[(unsigned long long)a.model_idx[kk] * a.eigen_stride + 2 * s * sp + sp + threadIdx.x];
        s_aux[16 + threadIdx.x] = t;
      }
      __syncthreads();
      dmma_build_b(s_diag, s_aux, s_bt, s, sp, K * sp);
      __syncthreads();
      dmma_pass(a.sum, a.weights, a.invariant, (any_pinv && a.invariant) ? s_aux + 16 : nullptr, a.sites, s_bt, K * sp,
                (unsigned long long)blockIdx.x * (NW_THREADS / 32) + warp,
                (unsigned long long)gridDim.x * (NW_THREADS / 32), fwd, lane, acc1, acc2);
    }
    else if (MODE == 2)
    {
      /* K = 4: a thread's 4 elements belong to 4 different sites and one rate.  Register slot j holds
       * trip it0 + (j ^ perm), perm = bit-reversed rate index, so that two exchange rounds with FIXED
       * send/keep slots (2,3 -> 0,1 then 1 -> 0) leave in slot 0 the sum over the 4 rates of a
       * different site in each lane: 9 shuffles per 4 elements, and the division runs on all lanes. */
      const unsigned int perm = ((k & 1u) << 1) | ((k >> 1) & 1u);
      const double * dgw = s_diag + NW_MAX_DIAG * 3 - 48 + (size_t)k * 12; /* diag * w_k (1 - pinv_k) */
      for (unsigned int it0 = 0; it0 < n_it; it0 += 4)
      {
        d4 v[4];
        double L0[4], L1[4], L2[4];
        const unsigned int itf = it0 + perm; /* slot 0's trip: the site this lane finishes */
        const unsigned int idxf = fwd ? itf : n_it - 1 - itf;
        const unsigned long long sitef = (base + (unsigned long long)idxf * stride) >> 2;
        const bool actf = itf < n_it && sitef < a.sites;
        unsigned int wgt = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
          const unsigned int it = it0 + ((unsigned int)j ^ perm);
          const unsigned int idx = fwd ? it : n_it - 1 - it;
          const unsigned long long e = base + (unsigned long long)idx * stride;
          v[j].x = v[j].y = v[j].z = v[j].w = 0.0;
          if (it < n_it && e < E) v[j] = ldg256(a.sum + e * 4);
        }
        if (actf) wgt = __ldg(a.weights + sitef);
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
          L0[j] = v[j].x * dgw[0] + v[j].y * dgw[3] + v[j].z * dgw[6] + v[j].w * dgw[9];
          L1[j] = v[j].x * dgw[1] + v[j].y * dgw[4] + v[j].z * dgw[7] + v[j].w * dgw[10];
          L2[j] = v[j].x * dgw[2] + v[j].y * dgw[5] + v[j].z * dgw[8] + v[j].w * dgw[11];
        }
        if (any_pinv && a.invariant)
        {
#pragma unroll
          for (int j = 0; j < 4; ++j)
          {
            const unsigned int it = it0 + ((unsigned int)j ^ perm);
            const unsigned int idx = fwd ? it : n_it - 1 - it;
            const unsigned long long e = base + (unsigned long long)idx * stride;
            if (it < n_it && e < E)
            {
              const int inv = a.invariant[e >> 2];
              if (inv >= 0) L0[j] += fr[inv] * pinv * wk;
            }
          }
        }
        L0[0] += __shfl_xor_sync(0xffffffffu, L0[2], 1);
        L0[1] += __shfl_xor_sync(0xffffffffu, L0[3], 1);
        L1[0] += __shfl_xor_sync(0xffffffffu, L1[2], 1);
        L1[1] += __shfl_xor_sync(0xffffffffu, L1[3], 1);
        L2[0] += __shfl_xor_sync(0xffffffffu, L2[2], 1);
        L2[1] += __shfl_xor_sync(0xffffffffu, L2[3], 1);
        L0[0] += __shfl_xor_sync(0xffffffffu, L0[1], 2);
        L1[0] += __shfl_xor_sync(0xffffffffu, L1[1], 2);
        L2[0] += __shfl_xor_sync(0xffffffffu, L2[1], 2);
        if (actf)
        {
          const double inv = 1.0 / L0[0]; /* one reciprocal for both ratios */
          const double r1 = -L1[0] * inv;
          acc1 += (double)wgt * r1;
          acc2 += (double)wgt * (r1 * r1 - L2[0] * inv);
        }
      }
    }
    else
    for (unsigned int it0 = 0; it0 < n_it; it0 += U)
    {
      bool act[U];
      d4 v[U];
      unsigned int wgt[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
      {
        const unsigned int it = it0 + u;
        const unsigned int idx = fwd ? it : n_it - 1 - it;
        const unsigned long long e = base + (unsigned long long)idx * stride;
        act[u] = it < n_it && e < E;
        wgt[u] = 0;
        if (act[u])
        {
          if (DNA) v[u] = ldg256(a.sum + e * 4);
          wgt[u] = __ldg(a.weights + (e >> kshift));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
      {
        if (it0 + u >= n_it) break; /* warp-uniform */
        const unsigned int idx = fwd ? it0 + u : n_it - 1 - (it0 + u);
        const unsigned long long e = base + (unsigned long long)idx * stride;
        const unsigned long long site = e >> kshift;
        double L0 = 0.0, L1 = 0.0, L2 = 0.0;
        if (act[u])
        {
          if (DNA)
          {
            L0 = v[u].x * dg[0] + v[u].y * dg[3] + v[u].z * dg[6] + v[u].w * dg[9];
            L1 = v[u].x * dg[1] + v[u].y * dg[4] + v[u].z * dg[7] + v[u].w * dg[10];
            L2 = v[u].x * dg[2] + v[u].y * dg[5] + v[u].z * dg[8] + v[u].w * dg[11];
          }
          else
          {
            const double * sm = a.sum + e * sp;
            if (s == 20)
            {
              /* protein: the element's 160 bytes as five 256-bit loads, all in flight before the first FMA */
              d4 q[5];
#pragma unroll
              for (int c = 0; c < 5; ++c) q[c] = ldg256(sm + 4 * c);
#pragma unroll
              for (int c = 0; c < 5; ++c)
              {
                const double * g = dg + 12 * c;
                L0 += q[c].x * g[0] + q[c].y * g[3] + q[c].z * g[6] + q[c].w * g[9];
                L1 += q[c].x * g[1] + q[c].y * g[4] + q[c].z * g[7] + q[c].w * g[10];
                L2 += q[c].x * g[2] + q[c].y * g[5] + q[c].z * g[8] + q[c].w * g[11];
              }
            }
            else
              for (unsigned int j = 0; j < s; ++j)
              {
                const double sv = sm[j];
                L0 += sv * dg[j * 3 + 0];
                L1 += sv * dg[j * 3 + 1];
                L2 += sv * dg[j * 3 + 2];
              }
          }
          if (pinv > 0.0)
          {
            double inv_lk = 0.0;
            if (a.invariant)
            {
              const int inv = a.invariant[site];
              if (inv >= 0) inv_lk = fr[inv] * pinv;
            }
            L0 = L0 * (1.0 - pinv) + inv_lk;
            L1 = L1 * (1.0 - pinv);
            L2 = L2 * (1.0 - pinv);
          }
          L0 *= wk;
          L1 *= wk;
          L2 *= wk;
        }
        for (unsigned int off = 1; off < K; off <<= 1)
        {
          L0 += __shfl_xor_sync(0xffffffffu, L0, off);
          L1 += __shfl_xor_sync(0xffffffffu, L1, off);
          L2 += __shfl_xor_sync(0xffffffffu, L2, off);
        }
        if (act[u] && k == 0)
        {
          const double r1 = -L1 / L0;
          acc1 += (double)wgt[u] * r1;
          acc2 += (double)wgt[u] * (r1 * r1 - L2 / L0);
        }
      }
    }
    /* CTA partial */
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
    {
      acc1 += __shfl_down_sync(0xffffffffu, acc1, off);
      acc2 += __shfl_down_sync(0xffffffffu, acc2, off);
    }
    if (lane == 0)
    {
      s_red[0][warp] = acc1;
      s_red[1][warp] = acc2;
    }
    __syncthreads();
    double * buf = a.partials + (size_t)(iter & 1u) * 2 * gridDim.x;
    if (threadIdx.x == 0)
    {
      double p1 = 0.0, p2 = 0.0;
      for (int wi = 0; wi < NW_THREADS / 32; ++wi)
      {
        p1 += s_red[0][wi];
        p2 += s_red[1][wi];
      }
      buf[blockIdx.x] = p1;
      buf[gridDim.x + blockIdx.x] = p2;
    }
    grid.sync();
    /* every CTA: the same ordered sum of all partials */
    {
      double t1, t2;
      newton_sum_partials(buf, gridDim.x, s_red, t1, t2);
      if (threadIdx.x == 0)
      {
        if (a.nranks > 1 && !newton_allgather(a, pass0 + iter, t1, t2)) t1 = t2 = nan("");
        s_tot[0] = t1;
        s_tot[1] = t2;
      }
    }
    __syncthreads();
    f = a.scale * s_tot[0];
    df = a.scale * a.scale * s_tot[1];

    /* damped Newton step inside the shrinking bracket (opt_algorithms.c:187-252) */
    if (!isfinite(f) || !isfinite(df))
    {
      status = NW_BAD_DERIV;
      break;
    }
    double dx;
    if (df > 0.0)
    {
      if (fabs(f) < a.tol) break;
      if (f < 0.0)
        xl = x;
      else
        xh = x;
      dx = -f / df;
    }
    else
      dx = -f / fabs(df);
    dx = fmax(fmin(dx, dxmax), -dxmax);
    if (x + dx < xl) dx = xl - x;
    if (x + dx > xh) dx = xh - x;
    if (fabs(dx) < a.tol) break;
    x += dx;
    x = fmax(fmin(x, a.xmax), a.xmin);
  }

  if (blockIdx.x == 0 && threadIdx.x == 0)
  {
    newton_publish(a, x, xorig, status, iter);
    if (a.nranks > 1) /* passes really made: the iteration-limit exit counted one it did not run */
      *reinterpret_cast<unsigned long long *>(a.inbox[a.rank] + NW_XCHG_COUNTER) =
          pass0 + iter - (status == NW_LIMIT ? 1u : 0u);
    a.result_dev[0] = x;
    a.result_dev[1] = status;
    a.result_dev[2] = (double)iter;
    a.result_dev[3] = f;
    a.result_host[0] = x;
    a.result_host[1] = status;
    a.result_host[2] = (double)iter;
    a.result_host[3] = f;
    __threadfence_system();
    *reinterpret_cast<volatile unsigned long long *>(a.tag_host) = a.tag;
  }
}

/* ---- 4 states, 4 rates: the same iteration with the sumtable streamed through a TMA ring ----------
 *
 * The register version above is latency-bound: a warp issues its four 32-byte loads, waits ~2 us under
 * load, computes, and only then asks for more (13 trips per warp and pass at 1 M sites; ncu: 25 % of
 * the warps resident, issue slots 29 % busy, DRAM 20-33 %).  Here the table arrives in 32 KB tiles
 * (1024 elements = 256 sites, plus their 256 pattern weights) by cp.async.bulk into a 3-stage shared
 * memory ring, completion on one mbarrier per stage: two stages (66 KB) per CTA are always in flight
 * without holding a register, one elected thread refills a stage as soon as the CTA has lifted it into
 * registers.  Thread t owns elements {j*256 + t} of a tile: one rate (t & 3) of four different sites,
 * reduced over the rates with the fixed-slot exchange described above.
 */
#define NT_TILE 1024 /* elements per tile */
#define NT_STAGES 3
#define NT_THREADS (NW_THREADS + 32) /* 8 consumer warps + 1 producer warp */

struct NewtonTmaSmem
{
  double tile[NT_STAGES][NT_TILE * 4];
  unsigned int wts[NT_STAGES][NT_TILE / 4];
  unsigned long long full[NT_STAGES];  /* TMA -> consumers: bytes have landed */
  unsigned long long empty[NT_STAGES]; /* consumers -> producer: stage is in registers (8 arrivals) */
  double diag[48]; /* [k][j][3], w_k (1 - pinv_k) folded in */
  double red[2][NW_THREADS / 32];
  double tot[2];
};

__device__ __forceinline__ void mbar_arrive(unsigned long long * bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <bool EVAL> /* EVAL: one pass, no grid barrier: an ordinary (non-cooperative) launch */
__global__ void __launch_bounds__(NT_THREADS, 2) newton_dna4_tma_kernel(NewtonArgs a)
{
  cg::grid_group grid = cg::this_grid();
  extern __shared__ __align__(128) unsigned char nt_smem_raw[];
  NewtonTmaSmem & sm = *reinterpret_cast<NewtonTmaSmem *>(nt_smem_raw);

  const unsigned int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const bool producer = warp == NW_THREADS / 32;
  const unsigned int k = tid & 3u;
  const unsigned int perm = ((k & 1u) << 1) | ((k >> 1) & 1u);
  const double wk = a.w[k], pinv = a.pinv[k];
  const bool any_pinv = a.any_pinv != 0;
  const double * fr = a.eigen + (unsigned long long)a.model_idx[k] * a.eigen_stride + 36;
  const unsigned long long E = (unsigned long long)a.sites * 4;
  const unsigned int ntiles = (unsigned int)((E + NT_TILE - 1) / NT_TILE);
  const unsigned int G = gridDim.x, b = blockIdx.x;
  const unsigned int nt = b < ntiles ? (ntiles - b + G - 1) / G : 0u; /* tiles of this CTA per pass */
  /* conflict-free 16-byte halves: lanes 4-7 of every 8 read the upper half first */
  const unsigned int swz = (tid >> 2) & 1u;

  if (tid == 0)
  {
    for (int st = 0; st < NT_STAGES; ++st)
    {
      mbar_init(&sm.full[st], 1);
      mbar_init(&sm.empty[st], NW_THREADS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double xl = a.xmin, xh = a.xmax;
  const double xorig = (a.brlen_io && a.x0_from_device) ? __ldcg(a.brlen_io) : a.x0;
  /* passes done by earlier launches (device-resident: the host does not know how many a launch needs) */
  const unsigned long long pass0 =
      a.nranks > 1 ? __ldcg(reinterpret_cast<const unsigned long long *>(a.inbox[a.rank] + NW_XCHG_COUNTER)) : 0ull;
  double x = fmax(fmin(xorig, a.xmax), a.xmin);
  const double dxmax = a.xmax / a.max_iters;
  unsigned int iter = 0;
  unsigned int used = 0; /* tiles through the ring so far (all passes): stage = n % 3, fill number n / 3 */
  unsigned int prefetched = 0; /* tiles of the coming pass the producer has already asked for */
  double status = NW_OK;
  double f = 0.0, df = 0.0;

  for (;;)
  {
    if (iter++ > a.max_iters)
    {
      status = NW_LIMIT;
      break;
    }
    const bool fwd = a.eval_only ? a.eval_only == 2 : (iter & 1u) == 0;
    double acc1 = 0.0, acc2 = 0.0;

    if (producer)
    {
      if (lane == 0)
      {
        /* ring position n (all passes), pass-local tile i in direction dir */
        auto issue = [&](unsigned int n, unsigned int i, bool dir) {
          const unsigned int slot = n % NT_STAGES, fill = n / NT_STAGES;
          if (fill) mbar_wait(&sm.empty[slot], (fill - 1u) & 1u); /* previous contents consumed */
          const unsigned int mine = dir ? i : nt - 1 - i;
          const unsigned long long e0 = ((unsigned long long)b + (unsigned long long)mine * G) * NT_TILE;
          const unsigned int elems = (unsigned int)((E - e0) < NT_TILE ? (E - e0) : NT_TILE);
          const unsigned int wbytes = ((elems / 4 + 3u) & ~3u) * 4u; /* weights are padded to whole tiles */
          mbar_expect_tx(&sm.full[slot], elems * 32u + wbytes);
          tma_load_1d(sm.tile[slot], a.sum + e0 * 4, elems * 32u, &sm.full[slot]);
          tma_load_1d(sm.wts[slot], a.weights + e0 / 4, wbytes, &sm.full[slot]);
        };
        for (unsigned int i = prefetched; i < nt; ++i) issue(used + i, i, fwd);
        /* the table does not change between passes: start the next pass (opposite direction) now, so
         * that its first tiles land while the grid reduces; if there is no next pass they are drained
         * before the kernel ends */
        if (!a.eval_only)
          for (unsigned int i = 0; i < NT_STAGES && i < nt; ++i) issue(used + nt + i, i, !fwd);
      }
    }
    else if (tid < 16)
    {
      /* diagonal table for the current branch length, while the first tiles are on their way */
      const double t = a.scale * x;
      const unsigned int kk = tid >> 2, j = tid & 3u;
      const double ev = a.eigen[(unsigned long long)a.model_idx[kk] * a.eigen_stride + 32 + j];
      const double ki = a.rate[kk] / (1.0 - a.pinv[kk]);
      const double e0 = exp(ev * ki * t);
      const double wf = a.w[kk] * (1.0 - a.pinv[kk]);
      sm.diag[tid * 3 + 0] = wf * e0;
      sm.diag[tid * 3 + 1] = wf * ev * ki * e0;
      sm.diag[tid * 3 + 2] = wf * ev * ki * ev * ki * e0;
    }
    /* consumers only (named barrier 1): the table is ready */
    if (!producer)
    {
      asm volatile("bar.sync 1, %0;" ::"n"(NW_THREADS) : "memory");
      const double * dgw = sm.diag + k * 12;
      for (unsigned int i = 0; i < nt; ++i)
      {
        const unsigned int n = used + i, slot = n % NT_STAGES;
        const unsigned int mine = fwd ? i : nt - 1 - i;
        const unsigned long long e0 = ((unsigned long long)b + (unsigned long long)mine * G) * NT_TILE;
        const unsigned int elems = (unsigned int)((E - e0) < NT_TILE ? (E - e0) : NT_TILE);
        mbar_wait(&sm.full[slot], (n / NT_STAGES) & 1u);

        /* lift this thread's four elements and its site's weight out of the stage */
        d4 v[4];
        const unsigned int sitef = perm * 64u + (tid >> 2); /* tile-local site this lane finishes */
        const bool actf = sitef * 4u < elems;
        const unsigned int wgt = sm.wts[slot][sitef];
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
          const unsigned int el = (((unsigned int)j ^ perm) << 8) + tid;
          const double * src = sm.tile[slot] + (size_t)el * 4;
          const d2 h0 = *reinterpret_cast<const d2 *>(src + 2 * swz);
          const d2 h1 = *reinterpret_cast<const d2 *>(src + 2 * (swz ^ 1u));
          const d2 lo = swz ? h1 : h0, hi = swz ? h0 : h1;
          v[j].x = lo.x;
          v[j].y = lo.y;
          v[j].z = hi.x;
          v[j].w = hi.y;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[slot]); /* this warp no longer needs the stage */
        if (elems != NT_TILE)
        {
          /* last, partial tile of the table: whatever lies behind it in the stage counts as zero */
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if ((((unsigned int)j ^ perm) << 8) + tid >= elems) v[j].x = v[j].y = v[j].z = v[j].w = 0.0;
        }

        double L0[4], L1[4], L2[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
          L0[j] = v[j].x * dgw[0] + v[j].y * dgw[3] + v[j].z * dgw[6] + v[j].w * dgw[9];
          L1[j] = v[j].x * dgw[1] + v[j].y * dgw[4] + v[j].z * dgw[7] + v[j].w * dgw[10];
          L2[j] = v[j].x * dgw[2] + v[j].y * dgw[5] + v[j].z * dgw[8] + v[j].w * dgw[11];
        }
        if (any_pinv && a.invariant)
        {
#pragma unroll
          for (int j = 0; j < 4; ++j)
          {
            const unsigned int el = (((unsigned int)j ^ perm) << 8) + tid;
            if (el < elems)
            {
              const int inv = a.invariant[(e0 + el) >> 2];
              if (inv >= 0) L0[j] += fr[inv] * pinv * wk;
            }
          }
        }
        L0[0] += __shfl_xor_sync(0xffffffffu, L0[2], 1);
        L0[1] += __shfl_xor_sync(0xffffffffu, L0[3], 1);
        L1[0] += __shfl_xor_sync(0xffffffffu, L1[2], 1);
        L1[1] += __shfl_xor_sync(0xffffffffu, L1[3], 1);
        L2[0] += __shfl_xor_sync(0xffffffffu, L2[2], 1);
        L2[1] += __shfl_xor_sync(0xffffffffu, L2[3], 1);
        L0[0] += __shfl_xor_sync(0xffffffffu, L0[1], 2);
        L1[0] += __shfl_xor_sync(0xffffffffu, L1[1], 2);
        L2[0] += __shfl_xor_sync(0xffffffffu, L2[1], 2);
        if (actf)
        {
          const double inv = 1.0 / L0[0];
          const double r1 = -L1[0] * inv;
          acc1 += (double)wgt * r1;
          acc2 += (double)wgt * (r1 * r1 - L2[0] * inv);
        }
      }
      /* warp partial */
      if (!EVAL)
      {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
        {
          acc1 += __shfl_down_sync(0xffffffffu, acc1, off);
          acc2 += __shfl_down_sync(0xffffffffu, acc2, off);
        }
        if (lane == 0)
        {
          sm.red[0][warp] = acc1;
          sm.red[1][warp] = acc2;
        }
      }
    }
    if (EVAL)
    {
      /* d_f = sum, dd_f = sum (scale 1): block -> grid by ticket, the last CTA publishes result and tag */
      double v[2] = {acc1, acc2};
      ReduceOut r;
      r.partials = a.partials;
      r.ticket = a.ticket;
      r.result_dev = a.result_dev;
      r.result_host = a.result_host;
      r.tag_host = a.tag_host;
      r.tag = a.tag;
      grid_reduce_finish<2>(v, r);
      return;
    }
    used += nt;
    prefetched = a.eval_only ? 0u : (nt < NT_STAGES ? nt : NT_STAGES);
    __syncthreads();

    /* CTA partial -> grid total (deterministic: fixed order in every CTA) */
    double * buf = a.partials + (size_t)(iter & 1u) * 2 * G;
    if (tid == 0)
    {
      double p1 = 0.0, p2 = 0.0;
      for (int wi = 0; wi < NW_THREADS / 32; ++wi)
      {
        p1 += sm.red[0][wi];
        p2 += sm.red[1][wi];
      }
      buf[b] = p1;
      buf[G + b] = p2;
    }
    grid.sync();
    {
      double t1, t2;
      newton_sum_partials(buf, G, sm.red, t1, t2);
      if (tid == 0)
      {
        if (a.nranks > 1 && !newton_allgather(a, pass0 + iter, t1, t2)) t1 = t2 = nan("");
        sm.tot[0] = t1;
        sm.tot[1] = t2;
      }
    }
    __syncthreads();
    f = a.scale * sm.tot[0];
    df = a.scale * a.scale * sm.tot[1];

    /* damped Newton step inside the shrinking bracket (opt_algorithms.c:187-252) */
    if (!isfinite(f) || !isfinite(df))
    {
      status = NW_BAD_DERIV;
      break;
    }
    double dx;
    if (df > 0.0)
    {
      if (fabs(f) < a.tol) break;
      if (f < 0.0)
        xl = x;
      else
        xh = x;
      dx = -f / df;
    }
    else
      dx = -f / fabs(df);
    dx = fmax(fmin(dx, dxmax), -dxmax);
    if (x + dx < xl) dx = xl - x;
    if (x + dx > xh) dx = xh - x;
    if (fabs(dx) < a.tol) break;
    x += dx;
    x = fmax(fmin(x, a.xmax), a.xmin);
  }

  /* tiles requested for a pass that will not happen must land before the CTA gives up its shared memory */
  for (unsigned int i = 0; i < prefetched; ++i)
    mbar_wait(&sm.full[(used + i) % NT_STAGES], ((used + i) / NT_STAGES) & 1u);

  if (b == 0 && tid == 0)
  {
    newton_publish(a, x, xorig, status, iter);
    if (a.nranks > 1) /* passes really made: the iteration-limit exit counted one it did not run */
      *reinterpret_cast<unsigned long long *>(a.inbox[a.rank] + NW_XCHG_COUNTER) =
          pass0 + iter - (status == NW_LIMIT ? 1u : 0u);
    a.result_dev[0] = x;
    a.result_dev[1] = status;
    a.result_dev[2] = (double)iter;
    a.result_dev[3] = f;
    a.result_host[0] = x;
    a.result_host[1] = status;
    a.result_host[2] = (double)iter;
    a.result_host[3] = f;
    __threadfence_system();
    *reinterpret_cast<volatile unsigned long long *>(a.tag_host) = a.tag;
  }
}

/* ---- many partitions, one shared branch length ------------------------------------------------------
 *
 * With linked (or proportionally scaled) branch lengths the Newton iteration needs, per pass, the sum of
 * the derivatives of ALL partitions (pll_optimize.c:1223-1287): per partition and pass one launch and one
 * host wait — 580 of the 930 launches per branch of a 100-partition analysis.  Here one cooperative grid
 * walks every partition's sumtable: the work is cut into items of 256 threads x U elements (U = 4 for 4
 * states, 1 otherwise), each CTA owns a contiguous run of items (so it switches partition — rebuilds the
 * diagonal table — once or twice per pass), per-thread sums carry the chain-rule factors s_p and s_p^2,
 * and the grid barrier / ordered sum / damped update are those of newton_kernel.  Partition descriptors
 * (NewtonArgs, model part only) live in a device array that the host re-uploads only when they change.
 */
struct NewtonMultiArgs
{
  const NewtonArgs * parts;
  const unsigned int * item_first; /* [nparts + 1] prefix sums of the items per partition */
  unsigned int nparts, total_items;
  double xmin, xmax, x0, tol;
  unsigned int max_iters;
  double * partials;
  double * result_dev;
  double * result_host;
  unsigned long long * tag_host;
  unsigned long long tag;
  /* site-sharded ranks: the lead partition's exchange buffers (see NewtonArgs) */
  int rank, nranks;
  double * inbox[NW_MAXRANKS];
};

template <bool DNA>
__device__ __forceinline__ void newton_multi_item(const NewtonArgs & a, const double * s_diag, unsigned int tile,
                                                  int lane, double wk, double pinv, const double * fr,
                                                  double & acc1, double & acc2)
{
  constexpr int U = DNA ? 4 : 1;
  const unsigned int K = a.K, s = a.states, sp = a.sp;
  const unsigned int k = (unsigned int)lane & (K - 1);
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long base = (unsigned long long)tile * (NW_THREADS * U) + threadIdx.x;
  const double * dg = s_diag + (size_t)k * s * 3;
  bool act[U];
  d4 v[U];
  unsigned int wgt[U];
#pragma unroll
  for (int u = 0; u < U; ++u)
  {
    const unsigned long long e = base + (unsigned long long)u * NW_THREADS;
    act[u] = e < E;
    wgt[u] = 0;
    if (act[u])
    {
      if (DNA) v[u] = ldg256(a.sum + e * 4);
      wgt[u] = __ldg(a.weights + (e >> a.kshift));
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u)
  {
    const unsigned long long e = base + (unsigned long long)u * NW_THREADS;
    double L0 = 0.0, L1 = 0.0, L2 = 0.0;
    if (act[u])
    {
      if (DNA)
      {
        L0 = v[u].x * dg[0] + v[u].y * dg[3] + v[u].z * dg[6] + v[u].w * dg[9];
        L1 = v[u].x * dg[1] + v[u].y * dg[4] + v[u].z * dg[7] + v[u].w * dg[10];
        L2 = v[u].x * dg[2] + v[u].y * dg[5] + v[u].z * dg[8] + v[u].w * dg[11];
      }
      else
      {
        const double * sm = a.sum + e * sp;
        if (s == 20)
        {
          d4 q[5];
#pragma unroll
          for (int c = 0; c < 5; ++c) q[c] = ldg256(sm + 4 * c);
#pragma unroll
          for (int c = 0; c < 5; ++c)
          {
            const double * g = dg + 12 * c;
            L0 += q[c].x * g[0] + q[c].y * g[3] + q[c].z * g[6] + q[c].w * g[9];
            L1 += q[c].x * g[1] + q[c].y * g[4] + q[c].z * g[7] + q[c].w * g[10];
            L2 += q[c].x * g[2] + q[c].y * g[5] + q[c].z * g[8] + q[c].w * g[11];
          }
        }
        else
          for (unsigned int j = 0; j < s; ++j)
          {
            const double sv = sm[j];
            L0 += sv * dg[j * 3 + 0];
            L1 += sv * dg[j * 3 + 1];
            L2 += sv * dg[j * 3 + 2];
          }
      }
      if (pinv > 0.0)
      {
        double inv_lk = 0.0;
        if (a.invariant)
        {
          const int inv = a.invariant[e >> a.kshift];
          if (inv >= 0) inv_lk = fr[inv] * pinv;
        }
        L0 = L0 * (1.0 - pinv) + inv_lk;
        L1 = L1 * (1.0 - pinv);
        L2 = L2 * (1.0 - pinv);
      }
      L0 *= wk;
      L1 *= wk;
      L2 *= wk;
    }
    for (unsigned int off = 1; off < K; off <<= 1)
    {
      L0 += __shfl_xor_sync(0xffffffffu, L0, off);
      L1 += __shfl_xor_sync(0xffffffffu, L1, off);
      L2 += __shfl_xor_sync(0xffffffffu, L2, off);
    }
    if (act[u] && k == 0)
    {
      const double r1 = -L1 / L0;
      acc1 += a.scale * (double)wgt[u] * r1;
      acc2 += a.scale * a.scale * (double)wgt[u] * (r1 * r1 - L2 / L0);
    }
  }
}

__global__ void __launch_bounds__(NW_THREADS, 3) newton_multi_kernel(NewtonMultiArgs m)
{
  cg::grid_group grid = cg::this_grid();
  __shared__ double s_diag[NW_MAX_DIAG * 3];
  __shared__ double s_red[2][NW_THREADS / 32];
  __shared__ double s_tot[2];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  /* this CTA's run of items and the partition its first item belongs to */
  const unsigned int lo = (unsigned int)((unsigned long long)m.total_items * blockIdx.x / gridDim.x);
  const unsigned int hi = (unsigned int)((unsigned long long)m.total_items * (blockIdx.x + 1) / gridDim.x);
  unsigned int p_lo = 0;
  {
    unsigned int l = 0, h = m.nparts; /* last p with item_first[p] <= lo */
    while (l + 1 < h)
    {
      const unsigned int mid = (l + h) >> 1;
      if (__ldg(m.item_first + mid) <= lo)
        l = mid;
      else
        h = mid;
    }
    p_lo = l;
  }

  double xl = m.xmin, xh = m.xmax;
  const double xorig = m.x0;
  const unsigned long long pass0 =
      m.nranks > 1 ? __ldcg(reinterpret_cast<const unsigned long long *>(m.inbox[m.rank] + NW_XCHG_COUNTER)) : 0ull;
  double x = fmax(fmin(xorig, m.xmax), m.xmin);
  const double dxmax = m.xmax / m.max_iters;
  unsigned int iter = 0;
  double status = NW_OK;
  double f = 0.0, df = 0.0;

  for (;;)
  {
    if (iter++ > m.max_iters)
    {
      status = NW_LIMIT;
      break;
    }
    double acc1 = 0.0, acc2 = 0.0;
    unsigned int p = p_lo, cur = 0xffffffffu;
    double wk = 0.0, pinv = 0.0;
    const double * fr = nullptr;
    for (unsigned int item = lo; item < hi; ++item)
    {
      while (item >= __ldg(m.item_first + p + 1)) ++p;
      const NewtonArgs & a = m.parts[p];
      if (p != cur)
      {
        cur = p;
        const unsigned int K = a.K, s = a.states, sp = a.sp;
        const double t = a.scale * x;
        __syncthreads();
        for (unsigned int i = threadIdx.x; i < K * s; i += blockDim.x)
        {
          const unsigned int kk = i / s, j = i % s;
          const double ev = a.eigen[(unsigned long long)a.model_idx[kk] * a.eigen_stride + 2 * s * sp + j];
          const double ki = a.rate[kk] / (1.0 - a.pinv[kk]);
          const double e0 = exp(ev * ki * t);
          s_diag[i * 3 + 0] = e0;
          s_diag[i * 3 + 1] = ev * ki * e0;
          s_diag[i * 3 + 2] = ev * ki * ev * ki * e0;
        }
        __syncthreads();
        const unsigned int k = (unsigned int)lane & (K - 1);
        wk = a.w[k];
        pinv = a.pinv[k];
        fr = a.eigen + (unsigned long long)a.model_idx[k] * a.eigen_stride + 2 * s * sp + sp;
      }
      const unsigned int tile = item - __ldg(m.item_first + p);
      if (a.states == 4)
        newton_multi_item<true>(a, s_diag, tile, lane, wk, pinv, fr, acc1, acc2);
      else
        newton_multi_item<false>(a, s_diag, tile, lane, wk, pinv, fr, acc1, acc2);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
    {
      acc1 += __shfl_down_sync(0xffffffffu, acc1, off);
      acc2 += __shfl_down_sync(0xffffffffu, acc2, off);
    }
    if (lane == 0)
    {
      s_red[0][warp] = acc1;
      s_red[1][warp] = acc2;
    }
    __syncthreads();
    double * buf = m.partials + (size_t)(iter & 1u) * 2 * gridDim.x;
    if (threadIdx.x == 0)
    {
      double p1 = 0.0, p2 = 0.0;
      for (int wi = 0; wi < NW_THREADS / 32; ++wi)
      {
        p1 += s_red[0][wi];
        p2 += s_red[1][wi];
      }
      buf[blockIdx.x] = p1;
      buf[gridDim.x + blockIdx.x] = p2;
    }
    grid.sync();
    {
      double t1, t2;
      newton_sum_partials(buf, gridDim.x, s_red, t1, t2);
      if (threadIdx.x == 0)
      {
        if (m.nranks > 1 && !newton_allgather(m, pass0 + iter, t1, t2)) t1 = t2 = nan("");
        s_tot[0] = t1;
        s_tot[1] = t2;
      }
    }
    __syncthreads();
    f = s_tot[0];
    df = s_tot[1];

    /* damped Newton step inside the shrinking bracket (opt_algorithms.c:187-252) */
    if (!isfinite(f) || !isfinite(df))
    {
      status = NW_BAD_DERIV;
      break;
    }
    double dx;
    if (df > 0.0)
    {
      if (fabs(f) < m.tol) break;
      if (f < 0.0)
        xl = x;
      else
        xh = x;
      dx = -f / df;
    }
    else
      dx = -f / fabs(df);
    dx = fmax(fmin(dx, dxmax), -dxmax);
    if (x + dx < xl) dx = xl - x;
    if (x + dx > xh) dx = xh - x;
    if (fabs(dx) < m.tol) break;
    x += dx;
    x = fmax(fmin(x, m.xmax), m.xmin);
  }

  if (blockIdx.x == 0 && threadIdx.x == 0)
  {
    if (m.nranks > 1)
      *reinterpret_cast<unsigned long long *>(m.inbox[m.rank] + NW_XCHG_COUNTER) =
          pass0 + iter - (status == NW_LIMIT ? 1u : 0u);
    m.result_dev[0] = x;
    m.result_dev[1] = status;
    m.result_dev[2] = (double)iter;
    m.result_dev[3] = f;
    m.result_host[0] = x;
    m.result_host[1] = status;
    m.result_host[2] = (double)iter;
    m.result_host[3] = f;
    __threadfence_system();
    *reinterpret_cast<volatile unsigned long long *>(m.tag_host) = m.tag;
  }
}

/* can the fused path serve this partition? */
int g_opt_fused_newton = 1;
int g_opt_queued_newton = 1;

bool newton_fused_supported(const PartCtx * c)
{
  const pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats;
  if (p->attributes & PLL_ATTRIB_AB_MASK) return false;
  if (K == 0 || K > NW_MAXK || (K & (K - 1)) != 0) return false;
  if ((size_t)K * p->states > NW_MAX_DIAG) return false;
  return g_opt_fused_newton != 0;
}

/* the model part of a partition's descriptor: sumtable, weights, eigen system, rates */
static bool fill_newton_model(PartCtx * c, const unsigned int * params_indices, const double * d_sumtable,
                              double scale, NewtonArgs & a)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats, s = p->states;
  if (!dev_sync_inputs(c) || !dev_sync_eigen(c, params_indices)) return false;
  memset(&a, 0, sizeof(a));
  a.sum = d_sumtable;
  a.weights = c->d_weights;
  a.eigen = c->d_eigen;
  a.eigen_stride = c->eigen_stride;
  a.sites = p->sites;
  a.states = s;
  a.sp = p->states_padded;
  a.K = K;
  for (a.kshift = 0; (1u << a.kshift) < K; ++a.kshift) {}
  bool any_pinv = false;
  for (unsigned int k = 0; k < K; ++k)
  {
    a.model_idx[k] = params_indices[k];
    a.w[k] = p->rate_weights[k];
    a.pinv[k] = p->prop_invar[params_indices[k]];
    a.rate[k] = p->rates[k];
    any_pinv |= a.pinv[k] > 0.0;
  }
  if (any_pinv)
  {
    if (!p->invariant)
    {
      pllb200_set_error(PLL_ERROR_INVAR_NONEFOUND, "prop_invar > 0 but no invariant site array.");
      return false;
    }
    if (!c->d_invariant)
    {
      c->invariant_dirty = true;
      if (!dev_sync_inputs(c)) return false;
    }
    a.invariant = c->d_invariant;
  }
  a.any_pinv = any_pinv;
  a.scale = scale;
  a.nranks = 1;
  return true;
}

bool launch_newton(PartCtx * c, const unsigned int * params_indices, const double * d_sumtable,
                   double xmin, double x0, double xmax, double tol, unsigned int max_iters,
                   double scale, int slot, double * brlen_io, int x0_from_device, int revert_on_limit, int eval_only)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats, s = p->states;
  NewtonArgs a;
  if (!fill_newton_model(c, params_indices, d_sumtable, scale, a)) return false;
  a.xmin = xmin;
  a.xmax = xmax;
  a.x0 = x0;
  a.tol = tol;
  a.max_iters = max_iters;
  a.partials = c->d_partials;
  a.result_dev = c->d_result + slot * 4;
  a.result_host = c->h_result_dev + slot * 4;
  dev_slot_tag(c, slot, &a.tag_host, &a.tag);
  a.rank = c->xchg_rank;
  a.nranks = c->xchg_nranks > 1 ? c->xchg_nranks : 1;
  for (int r = 0; r < a.nranks && r < NW_MAXRANKS; ++r) a.inbox[r] = c->xchg_peer[r];
  a.brlen_io = brlen_io;
  a.x0_from_device = x0_from_device;
  a.revert_on_limit = revert_on_limit;
  a.eval_only = eval_only;
  a.ticket = c->d_ticket;
  if (eval_only) a.nranks = 1; /* the caller's reduction callback adds the ranks' derivatives */

  /* cooperative grid: never more CTAs than can be co-resident, never more than there is work */
  static int per_sm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int mode = s == 4 ? (K == 4 ? 2 : 1) : 0;
  if (mode == 2 && g_opt_fused_newton == 1) mode = 3; /* fused_newton=2 keeps the register version */
  if (s == 20 && p->states_padded == 20 && K == 4 && g_opt_fused_newton != 3) mode = 4; /* fused_newton=3: generic kernel */
  if (mode == 0 && g_opt_fused_newton == 1 && g_opt_generic_dmma && s > 16 && s <= 64 && K <= 16 &&
      (K * p->states_padded) % 16 == 0 && K * p->states_padded <= NW_DMMA_MAXE)
    mode = 5;
  if (eval_only && mode != 3)
    return pllb200_cuda_ok(cudaErrorInvalidValue, "derivative pass on the TMA ring: wrong shape");
  const void * fn = mode == 5   ? (const void *)newton_kernel<5>
                    : mode == 4 ? (const void *)newton_kernel<4>
                    : mode == 3 ? (eval_only ? (const void *)newton_dna4_tma_kernel<true> : (const void *)newton_dna4_tma_kernel<false>)
                    : mode == 2 ? (const void *)newton_kernel<2>
                    : mode == 1 ? (const void *)newton_kernel<1>
                                : (const void *)newton_kernel<0>;
  const size_t dyn = mode == 3 ? sizeof(NewtonTmaSmem) + 128 : 0;
  const unsigned int threads = mode == 3 ? NT_THREADS : NW_THREADS;
  const int pi = eval_only ? 6 : mode;
  if (!per_sm[pi])
  {
    if (dyn && cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn) != cudaSuccess)
      return pllb200_cuda_ok(cudaGetLastError(), "newton kernel shared memory");
    int nb = 0;
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, (int)threads, dyn);
    per_sm[pi] = (e == cudaSuccess && nb > 0) ? nb : 1;
  }
  const unsigned long long E = (unsigned long long)a.sites * K;
  unsigned long long blocks = mode == 3 ? (E + NT_TILE - 1) / NT_TILE : (E + NW_THREADS - 1) / NW_THREADS;
  if (mode == 4 || mode == 5) blocks = (((unsigned long long)a.sites + 7) / 8 + NW_THREADS / 32 - 1) / (NW_THREADS / 32); /* a warp per 8-site tile */
  unsigned long long cap = (unsigned long long)c->sm_count * std::min(per_sm[pi], 4);
  if (cap > PLLB200_MAX_RED_BLOCKS * 3 / 4) cap = PLLB200_MAX_RED_BLOCKS * 3 / 4; /* partials: 2 buffers x 2 values */
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  void * kargs[] = {&a};
  const cudaError_t e = eval_only ? cudaLaunchKernel(fn, dim3((unsigned)blocks), dim3(threads), kargs, dyn, c->stream)
                                  : cudaLaunchCooperativeKernel(fn, dim3((unsigned)blocks), dim3(threads), kargs, dyn, c->stream);
  ++g_kernel_launches;
  return pllb200_cuda_ok(e, "newton kernel launch");
}

/* One shared branch length over `count` partitions (linked / scaled branch lengths): descriptors in a
 * device array owned by the first partition's context, launch on its stream after the other partitions'
 * streams (their sumtables) by event, result in one of its slots. */
bool launch_newton_multi(PartCtx ** cs, unsigned int count, unsigned int * const * params_indices,
                         double * const * d_sumtables, const double * scales, double xmin, double x0,
                         double xmax, double tol, unsigned int max_iters, int slot)
{
  PartCtx * c0 = cs[0];
  const size_t desc_bytes = sizeof(NewtonArgs) * count;
  const size_t first_bytes = sizeof(unsigned int) * (count + 1);
  const size_t total = desc_bytes + ((first_bytes + 15) & ~(size_t)15);
  std::vector<unsigned char> & img = c0->multi_shadow_next;
  img.assign(total, 0);
  NewtonArgs * parts = reinterpret_cast<NewtonArgs *>(img.data());
  unsigned int * first = reinterpret_cast<unsigned int *>(img.data() + desc_bytes);
  unsigned long long items = 0;
  for (unsigned int i = 0; i < count; ++i)
  {
    if (!fill_newton_model(cs[i], params_indices[i], d_sumtables[i], scales ? scales[i] : 1.0, parts[i]))
      return false;
    first[i] = (unsigned int)items;
    const unsigned long long E = (unsigned long long)parts[i].sites * parts[i].K;
    const unsigned long long per = (unsigned long long)NW_THREADS * (parts[i].states == 4 ? 4 : 1);
    items += (E + per - 1) / per;
    if (items > 0x7fffffffull)
    {
      pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Too many sites for the multi-partition Newton kernel.");
      return false;
    }
  }
  first[count] = (unsigned int)items;
  if (c0->multi_cap < total)
  {
    if (c0->d_multi) cudaFree(c0->d_multi);
    c0->d_multi = nullptr;
    c0->multi_cap = 0;
    c0->multi_shadow.clear();
    if (!pllb200_cuda_ok(cudaMalloc(&c0->d_multi, total), "multi-partition Newton descriptors")) return false;
    c0->multi_cap = total;
  }
  if (c0->multi_shadow.size() != total || memcmp(c0->multi_shadow.data(), img.data(), total) != 0)
  {
    /* pageable source: the runtime stages it before returning, the copy itself is ordered on the stream */
    if (!pllb200_cuda_ok(cudaMemcpyAsync(c0->d_multi, img.data(), total, cudaMemcpyHostToDevice, c0->stream),
                         "multi-partition Newton descriptors H2D"))
      return false;
    g_h2d_bytes += total;
    c0->multi_shadow = img;
  }
  /* the other partitions' sumtables (and input uploads) were queued on their own streams */
  for (unsigned int i = 1; i < count && !g_batch; ++i) /* in batch mode everything already runs on one stream */
  {
    PartCtx * c = cs[i];
    if (c->stream == c0->stream) continue;
    if (!c->ev_batch &&
        !pllb200_cuda_ok(cudaEventCreateWithFlags(&c->ev_batch, cudaEventDisableTiming), "batch event"))
      return false;
    if (!pllb200_cuda_ok(cudaEventRecord(c->ev_batch, c->stream), "batch event record")) return false;
    if (!pllb200_cuda_ok(cudaStreamWaitEvent(c0->stream, c->ev_batch, 0), "batch event wait")) return false;
  }
  NewtonMultiArgs m;
  memset(&m, 0, sizeof(m));
  m.parts = reinterpret_cast<const NewtonArgs *>(c0->d_multi);
  m.item_first = reinterpret_cast<const unsigned int *>(reinterpret_cast<const unsigned char *>(c0->d_multi) + desc_bytes);
  m.nparts = count;
  m.total_items = (unsigned int)items;
  m.xmin = xmin;
  m.xmax = xmax;
  m.x0 = x0;
  m.tol = tol;
  m.max_iters = max_iters;
  m.partials = c0->d_partials;
  m.result_dev = c0->d_result + slot * 4;
  m.result_host = c0->h_result_dev + slot * 4;
  dev_slot_tag(c0, slot, &m.tag_host, &m.tag);
  m.rank = c0->xchg_rank;
  m.nranks = c0->xchg_nranks > 1 ? c0->xchg_nranks : 1;
  for (int r = 0; r < m.nranks && r < NW_MAXRANKS; ++r) m.inbox[r] = c0->xchg_peer[r];

  static int per_sm = 0;
  if (!per_sm)
  {
    int nb = 0;
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, newton_multi_kernel, NW_THREADS, 0);
    per_sm = (e == cudaSuccess && nb > 0) ? nb : 1;
  }
  unsigned long long blocks = items;
  unsigned long long cap = (unsigned long long)c0->sm_count * std::min(per_sm, 3);
  if (cap > PLLB200_MAX_RED_BLOCKS * 3 / 4) cap = PLLB200_MAX_RED_BLOCKS * 3 / 4;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  void * kargs[] = {&m};
  const cudaError_t e = cudaLaunchCooperativeKernel((const void *)newton_multi_kernel, dim3((unsigned)blocks),
                                                      dim3(NW_THREADS), kargs, 0, c0->stream);
  ++g_kernel_launches;
  return pllb200_cuda_ok(e, "multi-partition newton kernel launch");
}
