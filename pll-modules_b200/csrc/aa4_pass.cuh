/*
 * aa4_pass.cuh — one pass over a 20-state, 4-category sumtable (640 B per site) for the derivative sums
 * (core_derivatives.c:643-772), shared by the Newton kernel (k_newton.cu, MODE 4) and the stand-alone derivative
 * kernel (k_derivatives.cu).
 *
 * The table is a flat array of 32-byte items (4 states of one rate).  A warp takes tiles of 160 items (8 sites)
 * with five fully coalesced 1 KB loads; lane l owns items l, l+32, ..., l+128 of every tile, i.e. always the same
 * five (rate, state chunk) pairs, whose 12 factors each (w_k (1 - pinv_k) folded in) stay in registers for the
 * whole pass.  Item sums go through a per-warp shared-memory tile; 24 lanes add the 20 items of
 * (site, L | L' | L''), every third lane finishes a site.  [Round 1: one 160-byte element per thread = 32 lines per
 * warp load and a shared-memory read per FMA: L1TEX-bound at 3.2 TB/s (Newton) / 1.8 TB/s (derivatives).]
 */
#ifndef PLLB200_AA4_PASS_CUH_
#define PLLB200_AA4_PASS_CUH_

#include "devutil.cuh"

#define AA4_TILE 160 /* items per warp and trip */

/* diag: [rate k][state j][3] = {e, lambda k e, (lambda k)^2 e}; wf[k] = w_k (1 - pinv_k) */
__device__ __forceinline__ void aa4_load_factors(const double * diag, const double * wf, int lane, double (&fc)[5][4][3])
{
#pragma unroll
  for (int u = 0; u < 5; ++u)
  {
    const unsigned int pi = (unsigned int)(u * 32 + lane) % 20u;
    const unsigned int kk = pi / 5u, cc = pi % 5u;
    const double w = wf[kk];
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
      const double * g = diag + (size_t)(kk * 20u + cc * 4u + j) * 3;
      fc[u][j][0] = w * g[0];
      fc[u][j][1] = w * g[1];
      fc[u][j][2] = w * g[2];
    }
  }
}

/* tr: AA4_TILE * 3 doubles of shared memory private to the warp.  inv_lk: null, or per state i the invariant-site
 * term sum_k w_k pinv_k pi_k[i].  gw / nw: this warp's index and the number of warps walking the table.
 * acc1 += weight * (-L'/L), acc2 += weight * ((L'/L)^2 - L''/L) in the finishing lanes. */
__device__ __forceinline__ void aa4_pass(const double * __restrict__ sum, const unsigned int * __restrict__ weights,
                                         const int * __restrict__ invariant, const double * inv_lk,
                                         unsigned int sites, const double (&fc)[5][4][3], double * tr,
                                         unsigned long long gw, unsigned long long nw, bool fwd, int lane,
                                         double & acc1, double & acc2)
{
  const unsigned long long ntiles = ((unsigned long long)sites + 7) / 8;
  const unsigned long long items = (unsigned long long)sites * 20;
  const unsigned long long mytiles = gw < ntiles ? (ntiles - gw + nw - 1) / nw : 0;
  const unsigned int fs = (unsigned int)lane / 3u, fcmp = (unsigned int)lane % 3u; /* lanes 0..23: (site, component) */
  for (unsigned long long ti = 0; ti < mytiles; ++ti)
  {
    const unsigned long long tile = gw + (fwd ? ti : mytiles - 1 - ti) * nw;
    const unsigned long long i0 = tile * AA4_TILE;
    d4 v[5];
#pragma unroll
    for (int u = 0; u < 5; ++u)
    {
      const unsigned long long it = i0 + (unsigned int)(u * 32 + lane);
      v[u] = d4{0.0, 0.0, 0.0, 0.0};
      if (it < items) v[u] = ldg256(sum + it * 4);
    }
    const unsigned long long site = tile * 8 + fs;
    const bool finisher = lane < 24 && fcmp == 0 && site < sites;
    unsigned int wgt = 0;
    int inv = -1;
    if (finisher)
    {
      wgt = __ldg(weights + site);
      if (inv_lk && invariant) inv = invariant[site];
    }
#pragma unroll
    for (int u = 0; u < 5; ++u)
    {
      double * o = tr + (size_t)(u * 32 + lane) * 3;
      o[0] = v[u].x * fc[u][0][0] + v[u].y * fc[u][1][0] + v[u].z * fc[u][2][0] + v[u].w * fc[u][3][0];
      o[1] = v[u].x * fc[u][0][1] + v[u].y * fc[u][1][1] + v[u].z * fc[u][2][1] + v[u].w * fc[u][3][1];
      o[2] = v[u].x * fc[u][0][2] + v[u].y * fc[u][1][2] + v[u].z * fc[u][2][2] + v[u].w * fc[u][3][2];
    }
    __syncwarp();
    double L = 0.0;
    if (lane < 24)
    {
      const double * q = tr + (size_t)fs * 60 + fcmp;
#pragma unroll
      for (int i = 0; i < 20; ++i) L += q[3 * i];
    }
    const double L1 = __shfl_down_sync(0xffffffffu, L, 1), L2 = __shfl_down_sync(0xffffffffu, L, 2);
    if (finisher)
    {
      double L0 = L;
      if (inv >= 0) L0 += inv_lk[inv];
      const double r = 1.0 / L0;
      const double r1 = -L1 * r;
      acc1 += (double)wgt * r1;
      acc2 += (double)wgt * (r1 * r1 - L2 * r);
    }
    __syncwarp(); /* the tile buffer is rewritten by the next trip */
  }
}

#endif
