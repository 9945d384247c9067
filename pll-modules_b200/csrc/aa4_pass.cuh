/*
 * aa4_pass.cuh — one pass over a 20-state, 4-category sumtable (640 B per site) for the derivative sums
 * (core_derivatives.c:643-772), shared by the Newton kernel (k_newton.cu, MODE 4) and the stand-alone derivative
 * kernel (k_derivatives.cu).
 *
 * (L, L', L'')[site] = sum_{e < 80} sum[site][e] * F[e][c],  F[e] = w_k (1 - pinv_k) (1, l, l^2) exp(l t)
 * is a [sites x 80] x [80 x 3] product: it runs on the FP64 tensor cores as mma.sync.m8n8k4 tiles with M = 8 sites,
 * N = 8 columns (3 used), K = 4 table entries per step, 20 steps per tile.
 *   A fragment (lane = 4 row + col): sum[site0 + row][4 ks + col] — one 8-byte load per lane and step, a quad of
 *     lanes reads one 32-byte sector, all 20 loads of a tile are issued before the first mma;
 *   B fragment: F[4 ks + col][n = lane / 4] (zero for n >= 3) from a 5 KB shared table [ks][lane] rebuilt once per
 *     pass (conflict-free 8-byte reads);
 *   D fragment: lane 4 row holds (L, L'), lane 4 row + 1 holds (L'', 0): one shuffle, then every fourth lane
 *     finishes its site (one reciprocal, pattern weight, invariant-site term).
 * 20 loads + 20 shared reads + 20 mma per 8 sites instead of 96 FMA + 27 shuffles + selects (~700 instructions,
 * issue-bound at 60 us per 128 MB pass: profiles/r02c_newton_aa_ncu_summary.txt).
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
    const unsigned int ks = i >> 5, l = i & 31u, n = l >> 2, e = 4u * ks + (l & 3u);
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
    const double * src = sum + site * 80 + ((unsigned int)lane & 3u);
    double av[20];
#pragma unroll
    for (int ks = 0; ks < 20; ++ks) av[ks] = act ? __ldg(src + 4 * ks) : 0.0;
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
    for (int ks = 0; ks < 20; ++ks) aa4_dmma(d0, d1, av[ks], btl[ks * 32]);
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
