/*
 * aa4_pass.cuh — one pass over a 20-state, 4-category sumtable (640 B per site) for the derivative sums
 * (core_derivatives.c:643-772), shared by the Newton kernel (k_newton.cu, MODE 4) and the stand-alone derivative
 * kernel (k_derivatives.cu).
 *
 * A site is 20 items of 32 bytes (4 states of one rate).  Lane p < 20 of a warp owns item p of EVERY site, i.e. one
 * (rate, state chunk) pair, whose 12 factors (w_k (1 - pinv_k) folded in) are 24 registers; lanes 20..31 carry
 * zeros.  A trip takes 8 sites: 8 independent 256-bit loads per lane (each warp load is one contiguous 640-byte
 * run), then per site three partial sums per lane.  Register slot j holds site j ^ P (P = lane bits 4,3,2), so that
 * three exchange rounds with FIXED send/keep slots (4..7 -> 0..3, 2,3 -> 0,1, 1 -> 0) followed by a butterfly over
 * the low two lane bits leave (L, L', L'') of site P in every lane: 27 shuffles of doubles per 8 sites, no shared
 * memory, and every fourth lane finishes a site.
 *
 * [Round 2, first version: 160-item tiles with five (rate, chunk) pairs per lane = 60 factors in registers:
 *  218 registers, 8 warps per SM, 5.5 k cycles per tile — 60 us per pass over 128 MB, no faster than the generic
 *  kernel (profiles/r02_newton_aa_ncu_summary.txt).  Round 1: one 160-byte element per thread = 32 lines per warp
 *  load and a shared-memory read per FMA.]
 */
#ifndef PLLB200_AA4_PASS_CUH_
#define PLLB200_AA4_PASS_CUH_

#include "devutil.cuh"

#define AA4_SITES 8 /* sites per warp and trip */

/* diag: [rate k][state j][3] = {e, lambda k e, (lambda k)^2 e}; wf[k] = w_k (1 - pinv_k) */
__device__ __forceinline__ void aa4_load_factors(const double * diag, const double * wf, int lane, double (&fc)[4][3])
{
  const unsigned int p = (unsigned int)lane < 20u ? (unsigned int)lane : 0u;
  const unsigned int kk = p / 5u, cc = p % 5u;
  const double w = (unsigned int)lane < 20u ? wf[kk] : 0.0;
#pragma unroll
  for (int j = 0; j < 4; ++j)
  {
    const double * g = diag + (size_t)(kk * 20u + cc * 4u + j) * 3;
    fc[j][0] = w * g[0];
    fc[j][1] = w * g[1];
    fc[j][2] = w * g[2];
  }
}

/* inv_lk: null, or per state i the invariant-site term sum_k w_k pinv_k pi_k[i].  gw / nw: this warp's index and the
 * number of warps walking the table.  acc1 += weight * (-L'/L), acc2 += weight * ((L'/L)^2 - L''/L) in the
 * finishing lanes. */
__device__ __forceinline__ void aa4_pass(const double * __restrict__ sum, const unsigned int * __restrict__ weights,
                                         const int * __restrict__ invariant, const double * inv_lk,
                                         unsigned int sites, const double (&fc)[4][3], unsigned long long gw,
                                         unsigned long long nw, bool fwd, int lane, double & acc1, double & acc2)
{
  const unsigned long long ntiles = ((unsigned long long)sites + AA4_SITES - 1) / AA4_SITES;
  const unsigned long long mytiles = gw < ntiles ? (ntiles - gw + nw - 1) / nw : 0;
  const unsigned int P = (((unsigned int)lane >> 4) & 1u) << 2 | (((unsigned int)lane >> 3) & 1u) << 1 |
                         (((unsigned int)lane >> 2) & 1u);
  const bool loader = (unsigned int)lane < 20u;
  for (unsigned long long ti = 0; ti < mytiles; ++ti)
  {
    const unsigned long long tile = gw + (fwd ? ti : mytiles - 1 - ti) * nw;
    const unsigned long long site0 = tile * AA4_SITES;
    const double * src = sum + (site0 * 20 + (unsigned int)lane) * 4;
    d4 v[AA4_SITES];
#pragma unroll
    for (int j = 0; j < AA4_SITES; ++j)
    {
      const unsigned int s = (unsigned int)j ^ P;
      v[j] = d4{0.0, 0.0, 0.0, 0.0};
      if (loader && site0 + s < sites) v[j] = ldg256(src + s * 80);
    }
    const unsigned long long sitef = site0 + P;
    const bool finisher = ((unsigned int)lane & 3u) == 0u && sitef < sites;
    unsigned int wgt = 0;
    int inv = -1;
    if (finisher)
    {
      wgt = __ldg(weights + sitef);
      if (inv_lk && invariant) inv = invariant[sitef];
    }
    double L[AA4_SITES][3];
#pragma unroll
    for (int j = 0; j < AA4_SITES; ++j)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        L[j][c] = v[j].x * fc[0][c] + v[j].y * fc[1][c] + v[j].z * fc[2][c] + v[j].w * fc[3][c];
#pragma unroll
    for (int c = 0; c < 3; ++c)
    {
#pragma unroll
      for (int j = 0; j < 4; ++j) L[j][c] += __shfl_xor_sync(0xffffffffu, L[j + 4][c], 16);
#pragma unroll
      for (int j = 0; j < 2; ++j) L[j][c] += __shfl_xor_sync(0xffffffffu, L[j + 2][c], 8);
      L[0][c] += __shfl_xor_sync(0xffffffffu, L[1][c], 4);
      L[0][c] += __shfl_xor_sync(0xffffffffu, L[0][c], 2);
      L[0][c] += __shfl_xor_sync(0xffffffffu, L[0][c], 1);
    }
    if (finisher)
    {
      double L0 = L[0][0];
      if (inv >= 0) L0 += inv_lk[inv];
      const double r = 1.0 / L0;
      const double r1 = -L[0][1] * r;
      acc1 += (double)wgt * r1;
      acc2 += (double)wgt * (r1 * r1 - L[0][2] * r);
    }
  }
}

#endif
