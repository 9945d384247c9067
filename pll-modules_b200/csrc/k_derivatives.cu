/*
 * k_derivatives.cu — sumtable construction and first/second derivatives of -logL w.r.t. one
 * branch length, the inner loop of Newton-Raphson branch-length optimisation.
 *
 * Replaces pll_core_update_sumtable_ii / _ti and pll_core_likelihood_derivatives
 * (libs/libpll/src/core_derivatives.c:116-640, 643-929 and SIMD twins), driven like
 * derivatives.c:24-324 and :333-418.
 *
 *   sum[n,k,j] = (sum_i P[n,k,i] pi_k[i] Vinv_k[i][j]) * (sum_i V_k[j][i] C[n,k,i])          (A.6)
 *     tip end: P[n,k,i] := bit i of the tip state, and the tip is always the "P" side
 *     (derivatives.c:66-78); per-rate scalers: min-normalised, capped at 4 (core_derivatives.c:418-440)
 *   diag[k,j]  = ( e^{l_j r_k t/(1-pinv_k)}, l r e, (l r)^2 e ) computed on the host per call (K*s exps)
 *   (L,L',L'')_n = sum_k w_k [ sum_j sum[n,k,j] diag[k,j] ] (+ invariant term on L)              (A.7)
 *   d_f  = sum_n wgt_n (-L'/L),  dd_f = sum_n wgt_n ((L'/L)^2 - L''/L)
 *
 * The sumtable (one CLV-sized buffer) never leaves HBM: the caller's host pointer is only a key.
 * Sumtable: 384 B/site (ii) / 257 (ti); derivative pass: 132 B/site (DNA Gamma4) -> a 1 M-site pass
 * is ~20 us of HBM time, so the kernel returns its two scalars through mapped pinned memory.
 */
#include "internal.h"
#include "devutil.cuh"
#include "reduce.cuh"
#include "aa4_pass.cuh"

#include <cmath>
#include <utility>

#define DV_MAXK 32
#define DV_THREADS 256
#define DVG_THREADS 128

struct SumArgs
{
  const double * pclv;         /* "parent" side (pi * Vinv); null if tip */
  const unsigned char * ptip;
  const double * cclv;         /* child side (V) — always an inner CLV */
  const unsigned int * pscaler;
  const unsigned int * cscaler;
  const double * eigen;
  const unsigned long long * tipmap;
  double * sum;
  unsigned long long eigen_stride;
  unsigned int sites, states, sp, K;
  int rate_scalers;
  int is4;
  unsigned int model_idx[DV_MAXK];
};

#define SUM_U 2    /* elements per thread per trip: 2 x (2 x 32 B) loads in flight */
#define SUM_LD 18  /* row stride of the staged eigen tables in doubles: rows of different rates sit in
                      different banks (16 would put all K rows on the same bank) */

template <int K, int TAG> /* TAG: the static shared tables are one symbol per kernel */
__device__ __forceinline__ void sumtable_dna_body(const SumArgs & a)
{
  __shared__ __align__(16) double s_w[K][SUM_LD];  /* pi_i * Vinv[i][j] */
  __shared__ __align__(16) double s_v[K][SUM_LD];  /* V[j][i] */
  for (int t = threadIdx.x; t < 16 * K; t += blockDim.x)
  {
    const int k = t / 16, q = t % 16;
    const double * blk = a.eigen + (unsigned long long)a.model_idx[k] * a.eigen_stride;
    s_v[k][q] = blk[q];
    s_w[k][q] = blk[36 + q / 4] * blk[16 + q];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int k = lane % K;
  const double * w = s_w[k];
  const double * v = s_v[k];
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long Eround = (E + 31) / 32 * 32;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long e0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; e0 < Eround;
       e0 += SUM_U * stride)
  {
    d4 p[SUM_U], c[SUM_U];
    unsigned int sc[SUM_U];
    bool act[SUM_U];
#pragma unroll
    for (int u = 0; u < SUM_U; ++u)
    {
      const unsigned long long e = e0 + u * stride;
      act[u] = e < E;
      sc[u] = 0;
      if (act[u])
      {
        if (a.ptip)
        {
          const unsigned int code = a.ptip[e / K];
          p[u].x = (code & 1) ? 1.0 : 0.0;
          p[u].y = (code & 2) ? 1.0 : 0.0;
          p[u].z = (code & 4) ? 1.0 : 0.0;
          p[u].w = (code & 8) ? 1.0 : 0.0;
        }
        else
          p[u] = ldg256(a.pclv + e * 4);
        c[u] = ldg256(a.cclv + e * 4);
        if (a.rate_scalers)
        {
          if (a.pscaler) sc[u] += a.pscaler[e];
          if (a.cscaler) sc[u] += a.cscaler[e];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < SUM_U; ++u)
    {
      const unsigned long long e = e0 + u * stride;
      if (e >= Eround) break; /* warp-uniform */
      d4 out = {0, 0, 0, 0};
      if (act[u])
      {
        const d4 pp = p[u], cc = c[u];
        const double l0 = pp.x * w[0] + pp.y * w[4] + pp.z * w[8] + pp.w * w[12];
        const double l1 = pp.x * w[1] + pp.y * w[5] + pp.z * w[9] + pp.w * w[13];
        const double l2 = pp.x * w[2] + pp.y * w[6] + pp.z * w[10] + pp.w * w[14];
        const double l3 = pp.x * w[3] + pp.y * w[7] + pp.z * w[11] + pp.w * w[15];
        const double r0 = v[0] * cc.x + v[1] * cc.y + v[2] * cc.z + v[3] * cc.w;
        const double r1 = v[4] * cc.x + v[5] * cc.y + v[6] * cc.z + v[7] * cc.w;
        const double r2 = v[8] * cc.x + v[9] * cc.y + v[10] * cc.z + v[11] * cc.w;
        const double r3 = v[12] * cc.x + v[13] * cc.y + v[14] * cc.z + v[15] * cc.w;
        out.x = l0 * r0;
        out.y = l1 * r1;
        out.z = l2 * r2;
        out.w = l3 * r3;
      }
      if (a.rate_scalers)
      {
        unsigned int mn = sc[u];
#pragma unroll
        for (int off = 1; off < K; off <<= 1)
        {
          const unsigned int o = __shfl_xor_sync(0xffffffffu, mn, off);
          mn = o < mn ? o : mn;
        }
        unsigned int diff = sc[u] - mn;
        if (diff > PLL_SCALE_RATE_MAXDIFF) diff = PLL_SCALE_RATE_MAXDIFF;
        double f = 1.0;
        for (unsigned int q = 0; q < diff; ++q) f *= SCALE_THRESHOLD_D;
        if (diff)
        {
          out.x *= f;
          out.y *= f;
          out.z *= f;
          out.w *= f;
        }
      }
      if (act[u]) stg256(a.sum + e * 4, out);
    }
  }
}

template <int K>
__global__ void __launch_bounds__(DV_THREADS, 4) sumtable_dna_kernel(SumArgs a)
{
  sumtable_dna_body<K, 0>(a);
}

/* the same for many partitions in one launch (blockIdx.y = partition): a branch of a 100-partition
 * analysis otherwise costs 100 launches here (optimize.cpp, batch mode) */
template <int K>
__global__ void __launch_bounds__(DV_THREADS, 4) sumtable_dna_batch_kernel(const SumArgs * __restrict__ arr)
{
  __shared__ SumArgs sa;
  for (unsigned int i = threadIdx.x; i < sizeof(SumArgs) / 4; i += blockDim.x)
    reinterpret_cast<unsigned int *>(&sa)[i] = reinterpret_cast<const unsigned int *>(arr + blockIdx.y)[i];
  __syncthreads();
  sumtable_dna_body<K, 1>(sa);
}

__global__ void __launch_bounds__(DVG_THREADS) sumtable_generic_kernel(SumArgs a)
{
  __shared__ unsigned int s_sc[DVG_THREADS];
  const unsigned int K = a.K, s = a.states, sp = a.sp;
  const unsigned int spb = DVG_THREADS / K;
  const unsigned int lsite = threadIdx.x / K, k = threadIdx.x % K;
  const unsigned long long site = (unsigned long long)blockIdx.x * spb + lsite;
  const bool act = lsite < spb && site < a.sites;
  const unsigned long long e = site * K + k;
  unsigned int sc = 0;
  if (act && a.rate_scalers)
  {
    if (a.pscaler) sc += a.pscaler[e];
    if (a.cscaler) sc += a.cscaler[e];
  }
  double f = 1.0;
  if (a.rate_scalers)
  {
    s_sc[threadIdx.x] = sc;
    __syncthreads();
    if (act)
    {
      unsigned int mn = sc;
      for (unsigned int q = 0; q < K; ++q)
        if (s_sc[lsite * K + q] < mn) mn = s_sc[lsite * K + q];
      unsigned int diff = sc - mn;
      if (diff > PLL_SCALE_RATE_MAXDIFF) diff = PLL_SCALE_RATE_MAXDIFF;
      for (unsigned int q = 0; q < diff; ++q) f *= SCALE_THRESHOLD_D;
    }
  }
  if (!act) return;
  const double * blk = a.eigen + (unsigned long long)a.model_idx[k] * a.eigen_stride;
  const double * ev = blk;
  const double * iev = blk + s * sp;
  const double * fr = blk + 2 * s * sp + sp;
  unsigned long long mask = 0;
  if (a.ptip)
  {
    const unsigned int code = a.ptip[site];
    mask = a.is4 ? (unsigned long long)code : a.tipmap[code];
  }
  const double * pv = a.pclv ? a.pclv + e * sp : nullptr;
  const double * cv = a.cclv + e * sp;
  double * dst = a.sum + e * sp;
  for (unsigned int j = 0; j < s; ++j)
  {
    double l = 0.0, r = 0.0;
    for (unsigned int i = 0; i < s; ++i)
    {
      const double pi = pv ? pv[i] : (double)((mask >> i) & 1ull);
      l += pi * fr[i] * iev[i * sp + j];
      r += ev[j * sp + i] * cv[i];
    }
    dst[j] = l * r * f;
  }
  for (unsigned int j = s; j < sp; ++j) dst[j] = 0.0;
}

/* ---- derivatives ----------------------------------------------------------------------------- */

struct DerivArgs
{
  const double * sum;
  const unsigned int * weights;
  const int * invariant;
  const double * eigen;
  const double * diag; /* generic path: staged [K][s][3] */
  unsigned long long eigen_stride;
  unsigned int sites, states, sp, K;
  unsigned int model_idx[DV_MAXK];
  double w[DV_MAXK];
  double pinv[DV_MAXK];
  ReduceOut red;
};

struct DiagDna
{
  double d[8][4][3];
};

/* 132 B per site and pass: a pure stream.  A thread keeps DV_U independent 256-bit loads (and their pattern
 * weights) in flight per trip — with one load per trip the kernel was latency-bound at 1.6 TB/s (round 1) —, the
 * grid is one resident wave, and every other call on the same table walks it backwards (`reverse`), so that what
 * the previous call touched last is still in the 126 MB L2 (the Newton kernel's trick, k_newton.cu). */
#define DV_U 4
template <int K>
__global__ void __launch_bounds__(DV_THREADS, 3) deriv_dna_kernel(DerivArgs a, DiagDna dg, int reverse)
{
  const int lane = threadIdx.x & 31;
  const int k = lane % K;
  const double wk = a.w[k], pinv = a.pinv[k];
  double d0[4], d1[4], d2[4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
  {
    d0[j] = dg.d[k][j][0];
    d1[j] = dg.d[k][j][1];
    d2[j] = dg.d[k][j][2];
  }
  const double * fr = a.eigen + (unsigned long long)a.model_idx[k] * a.eigen_stride + 36;
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long Eround = (E + 31) / 32 * 32;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long trips = (Eround + stride - 1) / stride; /* per thread, the last one may be empty */
  double acc1 = 0.0, acc2 = 0.0;
  for (unsigned long long t0 = 0; t0 < trips; t0 += DV_U)
  {
    d4 s[DV_U];
    unsigned int wgt[DV_U];
    unsigned long long ee[DV_U];
#pragma unroll
    for (int u = 0; u < DV_U; ++u)
    {
      const unsigned long long t = t0 + u < trips ? (reverse ? trips - 1 - (t0 + u) : t0 + u) : trips;
      ee[u] = tid + t * stride; /* >= Eround for the whole warp when there is nothing to do */
      s[u] = d4{0.0, 0.0, 0.0, 0.0};
      wgt[u] = 0;
      if (ee[u] < E)
      {
        s[u] = ldg256(a.sum + ee[u] * 4);
        if (k == 0) wgt[u] = __ldg(a.weights + ee[u] / K);
      }
    }
#pragma unroll
    for (int u = 0; u < DV_U; ++u)
    {
      if (ee[u] >= Eround) continue; /* warp-uniform: Eround, stride and the warp's first tid are multiples of 32 */
      const bool act = ee[u] < E;
      double L0 = s[u].x * d0[0] + s[u].y * d0[1] + s[u].z * d0[2] + s[u].w * d0[3];
      double L1 = s[u].x * d1[0] + s[u].y * d1[1] + s[u].z * d1[2] + s[u].w * d1[3];
      double L2 = s[u].x * d2[0] + s[u].y * d2[1] + s[u].z * d2[2] + s[u].w * d2[3];
      if (pinv > 0.0)
      {
        double inv_lk = 0.0;
        if (act && a.invariant)
        {
          const int inv = a.invariant[ee[u] / K];
          if (inv >= 0) inv_lk = fr[inv] * pinv;
        }
        L0 = L0 * (1.0 - pinv) + inv_lk;
        L1 = L1 * (1.0 - pinv);
        L2 = L2 * (1.0 - pinv);
      }
      L0 *= wk;
      L1 *= wk;
      L2 *= wk;
#pragma unroll
      for (int off = 1; off < K; off <<= 1)
      {
        L0 += __shfl_xor_sync(0xffffffffu, L0, off);
        L1 += __shfl_xor_sync(0xffffffffu, L1, off);
        L2 += __shfl_xor_sync(0xffffffffu, L2, off);
      }
      if (act && k == 0)
      {
        const double w = (double)wgt[u];
        const double r1 = -L1 / L0;
        const double r2 = r1 * r1 - L2 / L0;
        acc1 += w * r1;
        acc2 += w * r2;
      }
    }
  }
  double v[2] = {acc1, acc2};
  grid_reduce_finish<2>(v, a.red);
}

/* 4 categories without invariant sites (the hot shape): the tile mapping of logl_dna4_fast_kernel
 * (k_likelihood.cu).  A warp takes 2 tiles of 32 sites per trip = 8 independent 256-bit loads per thread; a
 * thread's 4 elements of a tile belong to 4 different sites, and two exchange rounds leave the three sums over the
 * rates of a DIFFERENT site in each lane, so that the two divisions of a site run on all 32 lanes instead of on a
 * quarter of them (deriv_dna_kernel<4>: 63 us per 1 M sites with DRAM at 26 %, issue slots full of divergent
 * DDIV sequences).  Same arithmetic per site as deriv_dna_kernel<4>. */
#define DVF_U 2
__global__ void __launch_bounds__(DV_THREADS, 2) deriv_dna4_fast_kernel(DerivArgs a, DiagDna dg, int reverse)
{
  const unsigned int lane = threadIdx.x & 31u, k = lane & 3u;
  const unsigned int perm = ((k & 1u) << 1) | (k >> 1);
  const double wk = a.w[k];
  double d0[4], d1[4], d2[4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
  {
    d0[j] = dg.d[k][j][0];
    d1[j] = dg.d[k][j][1];
    d2[j] = dg.d[k][j][2];
  }
  const unsigned long long tiles = ((unsigned long long)a.sites + 31) / 32;
  const unsigned long long groups = (tiles + DVF_U - 1) / DVF_U;
  const unsigned long long nwarps = (unsigned long long)gridDim.x * (DV_THREADS / 32);
  double acc1 = 0.0, acc2 = 0.0;
  for (unsigned long long g0 = (unsigned long long)blockIdx.x * (DV_THREADS / 32) + (threadIdx.x >> 5); g0 < groups;
       g0 += nwarps)
  {
    const unsigned long long g = reverse ? groups - 1 - g0 : g0;
    d4 v[DVF_U][4];
    unsigned int wgt[DVF_U];
    bool actf[DVF_U];
#pragma unroll
    for (int u = 0; u < DVF_U; ++u)
    {
      const unsigned long long tile = g * DVF_U + u;
      const unsigned long long site0 = tile * 32 + (lane >> 2);
      const double * sp = a.sum + (tile * 128 + lane) * 4;
#pragma unroll
      for (int j = 0; j < 4; ++j)
      {
        const unsigned int r = (unsigned int)j ^ perm;
        v[u][j] = d4{0.0, 0.0, 0.0, 0.0};
        if (site0 + 8 * r < a.sites) v[u][j] = ldg256(sp + r * 128);
      }
      actf[u] = site0 + 8 * perm < a.sites;
      wgt[u] = actf[u] ? __ldg(a.weights + site0 + 8 * perm) : 0u;
    }
#pragma unroll
    for (int u = 0; u < DVF_U; ++u)
    {
      double L0[4], L1[4], L2[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
      {
        const d4 s = v[u][j];
        L0[j] = (s.x * d0[0] + s.y * d0[1] + s.z * d0[2] + s.w * d0[3]) * wk;
        L1[j] = (s.x * d1[0] + s.y * d1[1] + s.z * d1[2] + s.w * d1[3]) * wk;
        L2[j] = (s.x * d2[0] + s.y * d2[1] + s.z * d2[2] + s.w * d2[3]) * wk;
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
      if (actf[u])
      {
        const double w = (double)wgt[u];
        const double r1 = -L1[0] / L0[0];
        const double r2 = r1 * r1 - L2[0] / L0[0];
        acc1 += w * r1;
        acc2 += w * r2;
      }
    }
  }
  double vv[2] = {acc1, acc2};
  grid_reduce_finish<2>(vv, a.red);
}

/* 20 states, 4 categories, unpadded: the tensor-core pass of aa4_pass.cuh (644 B per site); a.diag is the staged
 * [k][j][3] table of the generic path */
#define DVA_THREADS 256
__global__ void __launch_bounds__(DVA_THREADS, 3) deriv_aa4_kernel(DerivArgs a, int reverse)
{
  __shared__ double s_aux[24];
  __shared__ double s_bt[AA4_BT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bool any_pinv = false;
  for (unsigned int kk = 0; kk < 4; ++kk) any_pinv |= a.pinv[kk] > 0.0;
  if (threadIdx.x < 4) s_aux[threadIdx.x] = a.w[threadIdx.x] * (1.0 - a.pinv[threadIdx.x]);
  if (threadIdx.x < 20)
  {
    double t = 0.0;
    for (unsigned int kk = 0; kk < 4; ++kk)
      t += a.w[kk] * a.pinv[kk] * a.eigen[(unsigned long long)a.model_idx[kk] * a.eigen_stride + 2 * 20 * 20 + 20 + threadIdx.x];
    s_aux[4 + threadIdx.x] = t;
  }
  __syncthreads();
  aa4_build_b(a.diag, s_aux, s_bt);
  __syncthreads();
  double acc1 = 0.0, acc2 = 0.0;
  aa4_pass(a.sum, a.weights, a.invariant, (any_pinv && a.invariant) ? s_aux + 4 : nullptr, a.sites, s_bt,
           (unsigned long long)blockIdx.x * (DVA_THREADS / 32) + warp,
           (unsigned long long)gridDim.x * (DVA_THREADS / 32), reverse == 0, lane, acc1, acc2);
  double v[2] = {acc1, acc2};
  grid_reduce_finish<2>(v, a.red);
}

/* 17..64 states, K * states_padded a multiple of 16 and <= DVD_MAXE: dmma_pass of aa4_pass.cuh */
#define DVD_MAXE 256
__global__ void __launch_bounds__(DVA_THREADS, 2) deriv_dmma_kernel(DerivArgs a, int reverse)
{
  __shared__ double s_aux[80];
  __shared__ double s_bt[DVD_MAXE * 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned int K = a.K, s = a.states, sp = a.sp;
  bool any_pinv = false;
  for (unsigned int kk = 0; kk < K; ++kk) any_pinv |= a.pinv[kk] > 0.0;
  if (threadIdx.x < K) s_aux[threadIdx.x] = a.w[threadIdx.x] * (1.0 - a.pinv[threadIdx.x]);
  if (threadIdx.x < s)
  {
    double t = 0.0;
    for (unsigned int kk = 0; kk < K; ++kk)
      t += a.w[kk] * a.pinv[kk] * a.eigen[(unsigned long long)a.model_idx[kk] * a.eigen_stride + 2 * s * sp + sp + threadIdx.x];
    s_aux[16 + threadIdx.x] = t;
  }
  __syncthreads();
  dmma_build_b(a.diag, s_aux, s_bt, s, sp, K * sp);
  __syncthreads();
  double acc1 = 0.0, acc2 = 0.0;
  dmma_pass(a.sum, a.weights, a.invariant, (any_pinv && a.invariant) ? s_aux + 16 : nullptr, a.sites, s_bt, K * sp,
            (unsigned long long)blockIdx.x * (DVA_THREADS / 32) + warp,
            (unsigned long long)gridDim.x * (DVA_THREADS / 32), reverse == 0, lane, acc1, acc2);
  double v[2] = {acc1, acc2};
  grid_reduce_finish<2>(v, a.red);
}

__global__ void __launch_bounds__(DVG_THREADS) deriv_generic_kernel(DerivArgs a)
{
  __shared__ double s_l[3][DVG_THREADS];
  const unsigned int K = a.K, s = a.states, sp = a.sp;
  const unsigned int spb = DVG_THREADS / K;
  const unsigned int lsite = threadIdx.x / K, k = threadIdx.x % K;
  const unsigned long long nblk = ((unsigned long long)a.sites + spb - 1) / spb;
  double acc1 = 0.0, acc2 = 0.0;
  for (unsigned long long blk = blockIdx.x; blk < nblk; blk += gridDim.x)
  {
    const unsigned long long site = blk * spb + lsite;
    const bool act = lsite < spb && site < a.sites;
    double L0 = 0.0, L1 = 0.0, L2 = 0.0;
    if (act)
    {
      const double * sm = a.sum + (site * K + k) * sp;
      const double * dg = a.diag + (size_t)k * s * 3;
      for (unsigned int j = 0; j < s; ++j)
      {
        const double v = sm[j];
        L0 += v * dg[j * 3 + 0];
        L1 += v * dg[j * 3 + 1];
        L2 += v * dg[j * 3 + 2];
      }
      const double pinv = a.pinv[k];
      if (pinv > 0.0)
      {
        double inv_lk = 0.0;
        const int inv = a.invariant ? a.invariant[site] : -1;
        if (inv >= 0)
          inv_lk = a.eigen[(unsigned long long)a.model_idx[k] * a.eigen_stride + 2 * s * sp + sp + inv] * pinv;
        L0 = L0 * (1.0 - pinv) + inv_lk;
        L1 = L1 * (1.0 - pinv);
        L2 = L2 * (1.0 - pinv);
      }
      L0 *= a.w[k];
      L1 *= a.w[k];
      L2 *= a.w[k];
    }
    s_l[0][threadIdx.x] = L0;
    s_l[1][threadIdx.x] = L1;
    s_l[2][threadIdx.x] = L2;
    __syncthreads();
    if (act && k == 0)
    {
      double t0 = 0, t1 = 0, t2 = 0;
      for (unsigned int q = 0; q < K; ++q)
      {
        t0 += s_l[0][threadIdx.x + q];
        t1 += s_l[1][threadIdx.x + q];
        t2 += s_l[2][threadIdx.x + q];
      }
      const double wgt = (double)a.weights[site];
      const double r1 = -t1 / t0;
      const double r2 = r1 * r1 - t2 / t0;
      acc1 += wgt * r1;
      acc2 += wgt * r2;
    }
    __syncthreads();
  }
  double v[2] = {acc1, acc2};
  grid_reduce_finish<2>(v, a.red);
}

/* ---- ascertainment-bias terms of the derivatives (core_derivatives.c:852-924) ------------------
 * Lewis / Felsenstein: (A0,A1,A2) = sum over the `states` extra sites of (L,L',L'') * 2^(-256 c_n);
 *   Lewis        d_f  += W A1/(A0-1),            dd_f += W ((A0-1) A2 - A1^2)/(A0-1)^2
 *   Felsenstein  d_f  -= Winv A1/A0,             dd_f -= Winv (A2 A0 - A1^2)/A0^2
 * Stamatakis needs nothing here: the main kernel simply runs over sites + states.
 * One CTA, thread n = state n, same stream as the reducing kernel whose result slot it rewrites. */
#define ASC_THREADS 64

__global__ void __launch_bounds__(ASC_THREADS) asc_deriv_kernel(DerivArgs a, unsigned int real_sites,
                                                               const unsigned int * pscaler,
                                                               const unsigned int * cscaler,
                                                               int asc_type, unsigned int weight_sum)
{
  __shared__ double s_l[3][ASC_THREADS];
  const unsigned int K = a.K, s = a.states, sp = a.sp, n = threadIdx.x;
  if (n < s)
  {
    const unsigned long long site = (unsigned long long)real_sites + n;
    double t0 = 0.0, t1 = 0.0, t2 = 0.0;
    for (unsigned int k = 0; k < K; ++k)
    {
      const double * sm = a.sum + (site * K + k) * sp;
      const double * dg = a.diag + (size_t)k * s * 3;
      double L0 = 0.0, L1 = 0.0, L2 = 0.0;
      for (unsigned int j = 0; j < s; ++j)
      {
        const double v = sm[j];
        L0 += v * dg[j * 3 + 0];
        L1 += v * dg[j * 3 + 1];
        L2 += v * dg[j * 3 + 2];
      }
      t0 += L0 * a.w[k];
      t1 += L1 * a.w[k];
      t2 += L2 * a.w[k];
    }
    int sc = 0;
    if (pscaler) sc += (int)pscaler[site];
    if (cscaler) sc += (int)cscaler[site];
    s_l[0][n] = scalbn(t0, -256 * sc);
    s_l[1][n] = scalbn(t1, -256 * sc);
    s_l[2][n] = scalbn(t2, -256 * sc);
  }
  __syncthreads();
  if (n != 0) return;
  double A0 = 0.0, A1 = 0.0, A2 = 0.0;
  unsigned int sum_w_inv = 0;
  for (unsigned int q = 0; q < s; ++q)
  {
    A0 += s_l[0][q];
    A1 += s_l[1][q];
    A2 += s_l[2][q];
    sum_w_inv += a.weights[real_sites + q];
  }
  double df = a.red.result_dev[0], ddf = a.red.result_dev[1];
  if (asc_type == PLL_ATTRIB_AB_LEWIS)
  {
    df += (double)weight_sum * (A1 / (A0 - 1.0));
    ddf += (double)weight_sum * (((A0 - 1.0) * A2 - A1 * A1) / ((A0 - 1.0) * (A0 - 1.0)));
  }
  else
  {
    df -= (double)sum_w_inv * (A1 / A0);
    ddf -= (double)sum_w_inv * (((A2 * A0) - A1 * A1) / (A0 * A0));
  }
  a.red.result_dev[0] = df;
  a.red.result_dev[1] = ddf;
  a.red.result_host[0] = df;
  a.red.result_host[1] = ddf;
  publish_tag(a.red);
}

/* ---- host ------------------------------------------------------------------------------------ */

bool launch_sumtable(PartCtx * c, unsigned int parent_clv, unsigned int child_clv,
                     int parent_scaler, int child_scaler, const unsigned int * params_indices,
                     double * d_sumtable)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats;
  if (K > DV_MAXK)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "rate_cats > %d is not supported.", DV_MAXK);
    return false;
  }
  if (parent_clv >= p->nodes || child_clv >= p->nodes || parent_scaler >= (int)p->scale_buffers ||
      child_scaler >= (int)p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid index in pll_update_sumtable.");
    return false;
  }
  const bool ptip = is_tip_pattern(c, parent_clv), ctip = is_tip_pattern(c, child_clv);
  if (ptip && ctip)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "pll_update_sumtable() was called for the tip-tip case!");
    return false;
  }
  if (!dev_sync_inputs(c) || !dev_sync_eigen(c, params_indices)) return false;
  if (p->states == 20 && p->states_padded == 20 && !(p->attributes & PLL_ATTRIB_RATE_SCALERS) && K <= 8)
  {
    /* the tip (if any) takes the "parent" role (derivatives.c:66-78) */
    const unsigned int first = ctip ? child_clv : parent_clv, second = ctip ? parent_clv : child_clv;
    return launch_sumtable_aa(c, ptip || ctip, first, second, params_indices, d_sumtable);
  }
  if (generic_dmma_usable(c) && (!(ptip || ctip) || c->d_tipmap))
  {
    const unsigned int first = ctip ? child_clv : parent_clv, second = ctip ? parent_clv : child_clv;
    return launch_sumtable_dmma(c, ptip || ctip, first, second, params_indices, d_sumtable);
  }
  SumArgs a;
  memset(&a, 0, sizeof(a));
  if (ptip || ctip)
  {
    /* the tip takes the "parent" role, the inner node's scaler is the only one (derivatives.c:66-78) */
    const unsigned int tip = ptip ? parent_clv : child_clv;
    const unsigned int inner = ptip ? child_clv : parent_clv;
    a.ptip = c->d_tipchars + (size_t)tip * c->tip_stride;
    a.cclv = clv_dev(c, inner);
    a.pscaler = scaler_dev(c, ptip ? child_scaler : parent_scaler);
  }
  else
  {
    a.pclv = clv_dev(c, parent_clv);
    a.cclv = clv_dev(c, child_clv);
    a.pscaler = scaler_dev(c, parent_scaler);
    a.cscaler = scaler_dev(c, child_scaler);
  }
  a.eigen = c->d_eigen;
  a.eigen_stride = c->eigen_stride;
  a.tipmap = c->d_tipmap;
  a.sum = d_sumtable;
  a.sites = c->sites_alloc;
  a.states = p->states;
  a.sp = p->states_padded;
  a.K = K;
  a.rate_scalers = (p->attributes & PLL_ATTRIB_RATE_SCALERS) != 0;
  a.is4 = p->states == 4;
  for (unsigned int k = 0; k < K; ++k) a.model_idx[k] = params_indices[k];

  if (p->states == 4 && (K == 1 || K == 2 || K == 4 || K == 8))
  {
    const unsigned long long E = (unsigned long long)a.sites * K;
    unsigned long long blocks = (E + DV_THREADS - 1) / DV_THREADS;
    const unsigned long long cap = (unsigned long long)c->sm_count * 4; /* one resident wave (launch bounds) */
    if (blocks > cap) blocks = cap;
    if (batch_collects(BK_SUM_DNA4) && K == 4)
    {
      batch_collect(BK_SUM_DNA4, &a, sizeof(a), (unsigned)blocks, 0);
      return true;
    }
    switch (K)
    {
      case 1: sumtable_dna_kernel<1><<<(unsigned)blocks, DV_THREADS, 0, xstream(c)>>>(a); break;
      case 2: sumtable_dna_kernel<2><<<(unsigned)blocks, DV_THREADS, 0, xstream(c)>>>(a); break;
      case 4: sumtable_dna_kernel<4><<<(unsigned)blocks, DV_THREADS, 0, xstream(c)>>>(a); break;
      default: sumtable_dna_kernel<8><<<(unsigned)blocks, DV_THREADS, 0, xstream(c)>>>(a); break;
    }
  }
  else
  {
    const unsigned int spb = DVG_THREADS / K;
    const unsigned int blocks = (a.sites + spb - 1) / spb;
    sumtable_generic_kernel<<<blocks, DVG_THREADS, 0, xstream(c)>>>(a);
  }
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "sumtable kernel launch");
}

bool launch_derivatives(PartCtx * c, int parent_scaler, int child_scaler, double branch_length,
                        const unsigned int * params_indices, const double * d_sumtable, int slot)
{
  /* ratios L'/L cancel the scaling; only the asc-bias terms need the scalers */
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats, s = p->states, sp = p->states_padded;
  if (K > DV_MAXK)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "rate_cats > %d is not supported.", DV_MAXK);
    return false;
  }
  const int asc_type = (int)(p->attributes & PLL_ATTRIB_AB_MASK);
  if (asc_type && (s > ASC_THREADS || (p->attributes & PLL_ATTRIB_RATE_SCALERS)))
  {
    pllb200_set_error(PLL_ERROR_AB_NOSUPPORT,
                      "Ascertainment bias correction needs states <= %d and per-site scalers.", ASC_THREADS);
    return false;
  }
  if (parent_scaler >= (int)p->scale_buffers || child_scaler >= (int)p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid scaler index in derivative call.");
    return false;
  }
  /* 4 states x 4 rates without ascertainment correction: one pass of the Newton kernel's TMA ring (k_newton.cu) —
   * the table streams through shared memory with two stages in flight per CTA instead of through registers */
  if (s == 4 && p->sites >= 4096 && K == 4 && !asc_type && g_opt_fast_evals == 1 && g_opt_fused_newton == 1 &&
      newton_fused_supported(c))
  {
    eval_timing_begin(c->stream);
    const int reverse = (c->deriv_last_table == d_sumtable) ? (c->deriv_reverse ^= 1) : (c->deriv_reverse = 0);
    c->deriv_last_table = d_sumtable;
    if (!launch_newton(c, params_indices, d_sumtable, branch_length, branch_length, branch_length, 0.0, 1, 1.0, slot,
                       nullptr, 0, 0, reverse ? 2 : 1))
      return false;
    eval_timing_end(c->stream);
    return true;
  }
  if (!dev_sync_inputs(c) || !dev_sync_eigen(c, params_indices)) return false;
  DerivArgs a;
  memset(&a, 0, sizeof(a));
  a.sum = d_sumtable;
  a.weights = c->d_weights;
  a.eigen = c->d_eigen;
  a.eigen_stride = c->eigen_stride;
  /* Stamatakis: the extra per-state sites are ordinary weighted sites (core_derivatives.c:733-741) */
  a.sites = asc_type == PLL_ATTRIB_AB_STAMATAKIS ? c->sites_alloc : p->sites;
  a.states = s;
  a.sp = sp;
  a.K = K;
  bool any_pinv = false;
  for (unsigned int k = 0; k < K; ++k)
  {
    a.model_idx[k] = params_indices[k];
    a.w[k] = p->rate_weights[k];
    a.pinv[k] = p->prop_invar[params_indices[k]];
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
  a.red.partials = c->d_partials;
  a.red.ticket = c->d_ticket;
  a.red.result_dev = c->d_result + slot * 4;
  a.red.result_host = c->h_result_dev + slot * 4;

  /* diagonal table on the host: K*s exponentials (core_derivatives.c:757-772) */
  const bool dna = s == 4 && (K == 1 || K == 2 || K == 4 || K == 8);
  const bool asc_terms = asc_type == PLL_ATTRIB_AB_LEWIS || asc_type == PLL_ATTRIB_AB_FELSENSTEIN;
  /* completion tag: published by the kernel that touches the slot last */
  unsigned long long * tag_host = nullptr;
  unsigned long long tag = 0;
  dev_slot_tag(c, slot, &tag_host, &tag);
  a.red.tag = tag;
  a.red.tag_host = asc_terms ? nullptr : tag_host;
  DiagDna dg;
  std::vector<double> diag;
  if (!dna || asc_terms) diag.resize((size_t)K * s * 3);
  for (unsigned int k = 0; k < K; ++k)
  {
    const double * evals = p->eigenvals[params_indices[k]];
    const double ki = p->rates[k] / (1.0 - a.pinv[k]);
    for (unsigned int j = 0; j < s; ++j)
    {
      const double e0 = exp(evals[j] * ki * branch_length);
      const double e1 = evals[j] * ki * e0;
      const double e2 = evals[j] * ki * evals[j] * ki * e0;
      if (dna)
      {
        dg.d[k][j][0] = e0;
        dg.d[k][j][1] = e1;
        dg.d[k][j][2] = e2;
      }
      if (!dna || asc_terms)
      {
        diag[((size_t)k * s + j) * 3 + 0] = e0;
        diag[((size_t)k * s + j) * 3 + 1] = e1;
        diag[((size_t)k * s + j) * 3 + 2] = e2;
      }
    }
  }
  eval_timing_begin(c->stream);
  if (dna)
  {
    const unsigned long long E = (unsigned long long)a.sites * K;
    unsigned long long blocks = (E + DV_THREADS - 1) / DV_THREADS;
    const unsigned long long cap = (unsigned long long)c->sm_count * 3; /* one resident wave at 3 CTAs per SM */
    if (blocks > cap) blocks = cap;
    if (blocks > PLLB200_MAX_RED_BLOCKS) blocks = PLLB200_MAX_RED_BLOCKS;
    /* repeated passes over one table (Newton-Raphson through the C API): alternate the direction */
    const int reverse = (c->deriv_last_table == d_sumtable) ? (c->deriv_reverse ^= 1) : (c->deriv_reverse = 0);
    c->deriv_last_table = d_sumtable;
    if (K == 4 && !any_pinv && g_opt_fast_evals)
    {
      const unsigned long long groups = (((unsigned long long)a.sites + 31) / 32 + DVF_U - 1) / DVF_U;
      blocks = (groups + DV_THREADS / 32 - 1) / (DV_THREADS / 32);
      if (blocks > (unsigned long long)c->sm_count * 2) blocks = (unsigned long long)c->sm_count * 2;
      if (blocks > PLLB200_MAX_RED_BLOCKS) blocks = PLLB200_MAX_RED_BLOCKS;
      if (blocks < 1) blocks = 1;
      deriv_dna4_fast_kernel<<<(unsigned)blocks, DV_THREADS, 0, c->stream>>>(a, dg, reverse);
    }
    else
    switch (K)
    {
      case 1: deriv_dna_kernel<1><<<(unsigned)blocks, DV_THREADS, 0, c->stream>>>(a, dg, reverse); break;
      case 2: deriv_dna_kernel<2><<<(unsigned)blocks, DV_THREADS, 0, c->stream>>>(a, dg, reverse); break;
      case 4: deriv_dna_kernel<4><<<(unsigned)blocks, DV_THREADS, 0, c->stream>>>(a, dg, reverse); break;
      default: deriv_dna_kernel<8><<<(unsigned)blocks, DV_THREADS, 0, c->stream>>>(a, dg, reverse); break;
    }
  }
  else if (s == 20 && sp == 20 && K == 4 && g_opt_fused_newton != 3)
  {
    a.diag = (const double *)dev_stage(c, diag.data(), sizeof(double) * diag.size());
    if (!a.diag) return false;
    unsigned long long blocks = (((unsigned long long)a.sites + 7) / 8 + DVA_THREADS / 32 - 1) / (DVA_THREADS / 32);
    if (blocks > (unsigned long long)c->sm_count * 3) blocks = (unsigned long long)c->sm_count * 3;
    const int reverse = (c->deriv_last_table == d_sumtable) ? (c->deriv_reverse ^= 1) : (c->deriv_reverse = 0);
    c->deriv_last_table = d_sumtable;
    deriv_aa4_kernel<<<(unsigned)blocks, DVA_THREADS, 0, c->stream>>>(a, reverse);
  }
  else if (generic_dmma_usable(c) && g_opt_fused_newton == 1 && K <= 16 && (K * sp) % 16 == 0 && K * sp <= DVD_MAXE)
  {
    a.diag = (const double *)dev_stage(c, diag.data(), sizeof(double) * diag.size());
    if (!a.diag) return false;
    unsigned long long blocks = (((unsigned long long)a.sites + 7) / 8 + DVA_THREADS / 32 - 1) / (DVA_THREADS / 32);
    if (blocks > (unsigned long long)c->sm_count * 2) blocks = (unsigned long long)c->sm_count * 2;
    if (blocks < 1) blocks = 1;
    const int reverse = (c->deriv_last_table == d_sumtable) ? (c->deriv_reverse ^= 1) : (c->deriv_reverse = 0);
    c->deriv_last_table = d_sumtable;
    deriv_dmma_kernel<<<(unsigned)blocks, DVA_THREADS, 0, c->stream>>>(a, reverse);
  }
  else
  {
    a.diag = (const double *)dev_stage(c, diag.data(), sizeof(double) * diag.size());
    if (!a.diag) return false;
    const unsigned int spb = DVG_THREADS / K;
    unsigned long long blocks = ((unsigned long long)a.sites + spb - 1) / spb;
    const unsigned long long cap = (unsigned long long)c->sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks > PLLB200_MAX_RED_BLOCKS) blocks = PLLB200_MAX_RED_BLOCKS;
    deriv_generic_kernel<<<(unsigned)blocks, DVG_THREADS, 0, c->stream>>>(a);
  }
  eval_timing_end(c->stream);
  ++g_kernel_launches;
  if (!pllb200_cuda_ok(cudaGetLastError(), "derivative kernel launch")) return false;
  if (asc_terms)
  {
    a.red.tag_host = tag_host;
    if (dna)
    {
      a.diag = (const double *)dev_stage(c, diag.data(), sizeof(double) * diag.size());
      if (!a.diag) return false;
    }
    asc_deriv_kernel<<<1, ASC_THREADS, 0, c->stream>>>(a, p->sites, scaler_dev(c, parent_scaler),
                                                       scaler_dev(c, child_scaler), asc_type,
                                                       p->pattern_weight_sum);
    ++g_kernel_launches;
    return pllb200_cuda_ok(cudaGetLastError(), "asc derivative kernel launch");
  }
  return true;
}

bool flush_sum_dna4(PartCtx * lead, const void * d_args, const BatchKind & k)
{
  sumtable_dna_batch_kernel<4><<<dim3(batch_grid_x(lead, k, 4), k.count), DV_THREADS, 0, lead->stream>>>(
      static_cast<const SumArgs *>(d_args));
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "batched sumtable kernel launch");
}
