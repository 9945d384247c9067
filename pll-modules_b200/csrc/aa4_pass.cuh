/*
 * aa4_pass.cuh — one pass over a 20-state, 4-category sumtable (640 B per site) for the derivative sums
 * (core_derivatives.c:643-772), shared by the Newton kernel (k_newton.cu, MODE 4) and the stand-alone derivative
 * kernel (k_derivatives.cu).
 *
 * (L, L', L'')[site] = sum_{e < 80} sum[site][e] * F[e][c],  F[e] = w_k (1 - pinv_k) (1, l, l^2) exp(l t)
 * is a [sites x 80] x [80 x 3] product: it runs on the FP64 tensor cores as mma.sync.m8n8k4 tiles with M = 8 sites,
 * N = 8 columns (3 used), K = 4 table entries per step, 20 steps per tile.
 *   A fragment (lane = 4 row + col): the K dimension may be permuted freely as long as B follows, so lane (row, col)
 *     loads five 256-bit pieces sum[site0 + row][16 g + 4 col .. + 3], g = 0..4 — a quad of lanes reads one whole
 *     128-byte line of its site, a warp load touches 8 full lines — and k-step (g, j) takes component j of piece g:
 *     table entry e = 16 g + 4 col + j.  (With one 8-byte load per k-step, 20 per lane, the pass was bound by L1
 *     tag lookups: 240 us per Newton launch under ncu against 203 for the FMA version; profiles/README.md.)
 *   B fragment: F[16 g + 4 col + j][n = lane / 4] (zero for n >= 3) from a 5 KB shared table [ks = 4 g + j][lane]
 *     rebuilt once per pass (conflict-free 8-byte reads);
 *   D fragment: lane 4 row holds (L, L'), lane 4 row + 1 holds (L'', 0): one shuffle, then every fourth lane
 *     finishes its site (one reciprocal, pattern weight, invariant-site term).
 * 5 loads + 20 shared reads + 20 mma per 8 sites instead of 96 FMA + 27 shuffles + selects (~700 instructions:
 * profiles/r02c_newton_aa_ncu_summary.csv).  In a smoothing pass all three variants take the same ~28 us per pass:
 * the 128 MB table is just larger than the 126 MB L2 and the pass runs at what DRAM + L2 deliver (4.6 TB/s).
 *
 * [History.  Round 2 b: a lane per (rate, state chunk), 8 sites per trip in registers / through a TMA ring:
 *  350 instructions per 4 sites, issue-bound.  Round 2 a: 160-item tiles with 60 factors per lane: 218 registers,
 *  8 warps per SM.  Round 1: one 160-byte element per thread = 32 lines per warp load.]
 */
#ifndef PLLB200_AA4_PASS_CUH_
#define PLLB200_AA4_PASS_CUH_

#include "devutil.cuh"

#define AA4_SITES 8        /* sites per warp and trip */
#define AA4_BT (20 * 32)   /* doubles of the shared B table */

__device__ __forceinline__ void aa4_dmma(double & c0, double & c1, double a, double b)
{
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

/* All threads of the CTA; the caller synchronises before the pass.
 * diag: [rate k][state j][3] = {e, lambda k e, (lambda k)^2 e}; wf[k] = w_k (1 - pinv_k) */
__device__ __forceinline__ void aa4_build_b(const double * diag, const double * wf, double * bt)
{
  for (unsigned int i = threadIdx.x; i < AA4_BT; i += blockDim.x)
  {
    const unsigned int ks = i >> 5, l = i & 31u, n = l >> 2, e = 16u * (ks >> 2) + 4u * (l & 3u) + (ks & 3u);
    bt[i] = n < 3u ? wf[e / 20u] * diag[e * 3u + n] : 0.0;
  }
}

/* inv_lk: null, or per state i the invariant-site term sum_k w_k pinv_k pi_k[i].  gw / nw: this warp's index and the
 * number of warps walking the table.  acc1 += weight * (-L'/L), acc2 += weight * ((L'/L)^2 - L''/L) in the
 * finishing lanes. */
__device__ __forceinline__ void aa4_pass(const double * __restrict__ sum, const unsigned int * __restrict__ weights,
                                         const int * __restrict__ invariant, const double * inv_lk,
                                         unsigned int sites, const double * bt, unsigned long long gw,
                                         unsigned long long nw, bool fwd, int lane, double & acc1, double & acc2)
{
  const unsigned long long ntiles = ((unsigned long long)sites + AA4_SITES - 1) / AA4_SITES;
  const unsigned long long mytiles = gw < ntiles ? (ntiles - gw + nw - 1) / nw : 0;
  const double * btl = bt + lane;
  for (unsigned long long ti = 0; ti < mytiles; ++ti)
  {
    const unsigned long long tile = gw + (fwd ? ti : mytiles - 1 - ti) * nw;
    const unsigned long long site = tile * AA4_SITES + ((unsigned int)lane >> 2);
    const bool act = site < sites;
    const double * src = sum + site * 80 + 4u * ((unsigned int)lane & 3u);
    d4 xv[5];
#pragma unroll
    for (int g = 0; g < 5; ++g)
    {
      xv[g] = d4{0.0, 0.0, 0.0, 0.0};
      if (act) xv[g] = ldg256(src + 16 * g);
    }
    const bool finisher = ((unsigned int)lane & 3u) == 0u && act;
    unsigned int wgt = 0;
    int inv = -1;
    if (finisher)
    {
      wgt = __ldg(weights + site);
      if (inv_lk && invariant) inv = invariant[site];
    }
    double d0 = 0.0, d1 = 0.0;
#pragma unroll
    for (int g = 0; g < 5; ++g)
    {
      aa4_dmma(d0, d1, xv[g].x, btl[(4 * g + 0) * 32]);
      aa4_dmma(d0, d1, xv[g].y, btl[(4 * g + 1) * 32]);
      aa4_dmma(d0, d1, xv[g].z, btl[(4 * g + 2) * 32]);
      aa4_dmma(d0, d1, xv[g].w, btl[(4 * g + 3) * 32]);
    }
    const double L2 = __shfl_down_sync(0xffffffffu, d0, 1);
    if (finisher)
    {
      double L0 = d0;
      if (inv >= 0) L0 += inv_lk[inv];
      const double r = 1.0 / L0;
      const double r1 = -d1 * r;
      acc1 += (double)wgt * r1;
      acc2 += (double)wgt * (r1 * r1 - L2 * r);
    }
  }
}

/* ---- the same pass for any table row of E = K * states_padded doubles with E % 16 == 0 (17..64 states) ---------
 * bt: E / 4 * 32 doubles.  diag: [k][j < s][3]; wf[k] = w_k (1 - pinv_k).  Padding entries of a table row are zero
 * (both sumtable kernels write them) and get factor zero. */
__device__ __forceinline__ void dmma_build_b(const double * diag, const double * wf, double * bt, unsigned int s,
                                             unsigned int sp, unsigned int E)
{
  for (unsigned int i = threadIdx.x; i < E * 8u; i += blockDim.x)
  {
    const unsigned int ks = i >> 5, l = i & 31u, n = l >> 2, e = 16u * (ks >> 2) + 4u * (l & 3u) + (ks & 3u);
    const unsigned int k = e / sp, j = e % sp;
    bt[i] = (n < 3u && j < s) ? wf[k] * diag[(k * s + j) * 3u + n] : 0.0;
  }
}

/* inv_lk: null, or per state the invariant-site term */
__device__ __forceinline__ void dmma_pass(const double * __restrict__ sum, const unsigned int * __restrict__ weights,
                                          const int * __restrict__ invariant, const double * inv_lk,
                                          unsigned int sites, const double * bt, unsigned int E, unsigned long long gw,
                                          unsigned long long nw, bool fwd, int lane, double & acc1, double & acc2)
{
  const unsigned long long ntiles = ((unsigned long long)sites + 7) / 8;
  const unsigned long long mytiles = gw < ntiles ? (ntiles - gw + nw - 1) / nw : 0;
  const unsigned int pieces = E / 16u;
  const double * btl = bt + lane;
  for (unsigned long long ti = 0; ti < mytiles; ++ti)
  {
    const unsigned long long tile = gw + (fwd ? ti : mytiles - 1 - ti) * nw;
    const unsigned long long site = tile * 8 + ((unsigned int)lane >> 2);
    const bool act = site < sites;
    const double * src = sum + site * E + 4u * ((unsigned int)lane & 3u);
    const bool finisher = ((unsigned int)lane & 3u) == 0u && act;
    unsigned int wgt = 0;
    int inv = -1;
    if (finisher)
    {
      wgt = __ldg(weights + site);
      if (inv_lk && invariant) inv = invariant[site];
    }
    double d0 = 0.0, d1 = 0.0;
    for (unsigned int g0 = 0; g0 < pieces; g0 += 8) /* 8 pieces = 64 registers in flight per lane */
    {
      d4 xv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
      {
        xv[u] = d4{0.0, 0.0, 0.0, 0.0};
        if (act && g0 + u < pieces) xv[u] = ldg256(src + 16 * (g0 + u));
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
      {
        if (g0 + u >= pieces) break; /* warp-uniform */
        const double * b = btl + (size_t)(4 * (g0 + u)) * 32;
        aa4_dmma(d0, d1, xv[u].x, b[0]);
        aa4_dmma(d0, d1, xv[u].y, b[32]);
        aa4_dmma(d0, d1, xv[u].z, b[64]);
        aa4_dmma(d0, d1, xv[u].w, b[96]);
      }
    }
    const double L2 = __shfl_down_sync(0xffffffffu, d0, 1);
    if (finisher)
    {
      double L0 = d0;
      if (inv >= 0) L0 += inv_lk[inv];
      const double r = 1.0 / L0;
      const double r1 = -d1 * r;
      acc1 += (double)wgt * r1;
      acc2 += (double)wgt * (r1 * r1 - L2 * r);
    }
  }
}

#endif
