/*
 * k_ancestral.cu — marginal ancestral state probabilities at an inner node.
 *
 * Replaces pll_compute_node_ancestral / _extbuf (libs/libpll/src/likelihood.c:639-823).  The reference
 * builds a temporary CLV  T[n,k,j] = N[n,k,j] * sum_i P[k][j][i] O[n,k,i]  with pll_core_update_partial_*
 * (identity matrix on the node's side), then per site
 *     a[n,j] = sum_k w_k pi_k[j] T[n,k,j],   a[n,:] /= sum_j a[n,j].
 * Here that is one pass, thread = (site, state): no temporary CLV, P and pi·w from shared memory when they
 * fit, result written as [site][state] doubles into a device buffer that the API copies to the caller.
 * A per-site scaling factor is common to all states of a site and cancels in the normalisation, so
 * per-site scalers are not read at all.  With PLL_ATTRIB_RATE_SCALERS the categories of one site may
 * carry different powers of 2^256; they are brought to the site's smallest count before the sum (the
 * reference adds them as stored, likelihood.c:733-741 — see DESIGN.md).
 * Algorithmic bytes per site (DNA G4): 2 CLVs in, 32 B out = 288 B; not on the hot path (one call per
 * inner node of a finished search).
 */
#include "internal.h"
#include "devutil.cuh"

#define AN_THREADS 128
#define AN_MAXK 32
#define AN_SMEM_DOUBLES 4096 /* P-matrix of K * s * sp doubles is staged when it fits */

struct AncArgs
{
  const double * node_clv;
  const double * other_clv;          /* null when the other end is a tip pattern */
  const unsigned char * other_tip;
  const unsigned int * node_scaler;  /* only read with rate scalers */
  const unsigned int * other_scaler;
  const double * pmat;               /* [K][s][sp] of the connecting edge */
  const double * eigen;
  const unsigned long long * tipmap;
  double * out;                      /* [sites][states] */
  unsigned long long eigen_stride;
  unsigned int sites, states, sp, K, group; /* group = threads per site = pow2 >= states (<= 32) */
  int rate_scalers, is4;
  unsigned int freq_idx[AN_MAXK];
  double w[AN_MAXK];
};

__global__ void __launch_bounds__(AN_THREADS) ancestral_kernel(AncArgs a)
{
  __shared__ double s_pm[AN_SMEM_DOUBLES];
  const unsigned int s = a.states, sp = a.sp, K = a.K;
  const bool staged = (size_t)K * s * sp <= AN_SMEM_DOUBLES;
  if (staged)
    for (unsigned int i = threadIdx.x; i < K * s * sp; i += blockDim.x) s_pm[i] = a.pmat[i];
  __syncthreads();
  const double * pm = staged ? s_pm : a.pmat;

  const unsigned int spb = AN_THREADS / a.group; /* sites per CTA and trip */
  const unsigned int j = threadIdx.x % a.group, lsite = threadIdx.x / a.group;
  const unsigned long long trips = ((unsigned long long)a.sites + spb - 1) / spb;
  for (unsigned long long t = blockIdx.x; t < trips; t += gridDim.x)
  {
    const unsigned long long n = t * spb + lsite;
    const bool act = n < a.sites && j < s;
    double acc = 0.0;
    if (act)
    {
      unsigned long long mask = 0;
      if (a.other_tip)
      {
        const unsigned int code = a.other_tip[n];
        mask = a.is4 ? (unsigned long long)code : a.tipmap[code];
      }
      unsigned int smin = 0;
      if (a.rate_scalers)
      {
        smin = 0xffffffffu;
        for (unsigned int k = 0; k < K; ++k)
        {
          unsigned int sc = 0;
          if (a.node_scaler) sc += a.node_scaler[n * K + k];
          if (a.other_scaler) sc += a.other_scaler[n * K + k];
          smin = sc < smin ? sc : smin;
        }
      }
      for (unsigned int k = 0; k < K; ++k)
      {
        const double * row = pm + ((size_t)k * s + j) * sp;
        double x = 0.0;
        if (a.other_clv)
        {
          const double * o = a.other_clv + (n * K + k) * sp;
          for (unsigned int i = 0; i < s; ++i) x += row[i] * o[i];
        }
        else
          for (unsigned int i = 0; i < s; ++i)
            if ((mask >> i) & 1ull) x += row[i];
        const double pi = a.eigen[(unsigned long long)a.freq_idx[k] * a.eigen_stride + 2 * s * sp + sp + j];
        double term = a.w[k] * pi * a.node_clv[(n * K + k) * sp + j] * x;
        if (a.rate_scalers)
        {
          unsigned int sc = 0;
          if (a.node_scaler) sc += a.node_scaler[n * K + k];
          if (a.other_scaler) sc += a.other_scaler[n * K + k];
          if (sc != smin) term = scalbn(term, -256 * (int)min(sc - smin, 8u));
        }
        acc += term;
      }
    }
    /* normalise over the states of the site: the group is a power of two inside one warp */
    double sum = acc;
    for (unsigned int off = 1; off < a.group; off <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
    if (act) a.out[n * s + j] = acc / sum;
  }
}

bool launch_ancestral(PartCtx * c, unsigned int node_clv, int node_scaler, unsigned int other_clv,
                      int other_scaler, unsigned int matrix_index, const unsigned int * freqs_indices,
                      double * d_out)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats, s = p->states;
  if (K > AN_MAXK || s > 32)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID,
                      "Ancestral states need rate_cats <= %d and states <= 32.", AN_MAXK);
    return false;
  }
  if (node_clv >= p->nodes || other_clv >= p->nodes || matrix_index >= p->prob_matrices ||
      node_scaler >= (int)p->scale_buffers || other_scaler >= (int)p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid index in pll_compute_node_ancestral.");
    return false;
  }
  if (is_tip_pattern(c, node_clv))
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Ancestral states are defined for inner nodes only.");
    return false;
  }
  if (!dev_sync_inputs(c) || !dev_sync_eigen(c, freqs_indices)) return false;
  AncArgs a;
  memset(&a, 0, sizeof(a));
  a.node_clv = clv_dev(c, node_clv);
  if (is_tip_pattern(c, other_clv))
    a.other_tip = c->d_tipchars + (size_t)other_clv * c->tip_stride;
  else
    a.other_clv = clv_dev(c, other_clv);
  a.rate_scalers = (p->attributes & PLL_ATTRIB_RATE_SCALERS) != 0;
  if (a.rate_scalers)
  {
    a.node_scaler = scaler_dev(c, node_scaler);
    a.other_scaler = a.other_tip ? nullptr : scaler_dev(c, other_scaler);
  }
  a.pmat = c->d_pmat + (size_t)matrix_index * c->pmat_elems;
  a.eigen = c->d_eigen;
  a.eigen_stride = c->eigen_stride;
  a.tipmap = c->d_tipmap;
  a.out = d_out;
  a.sites = p->sites;
  a.states = s;
  a.sp = p->states_padded;
  a.K = K;
  a.is4 = s == 4;
  for (a.group = 1; a.group < s; a.group <<= 1) {}
  for (unsigned int k = 0; k < K; ++k)
  {
    a.freq_idx[k] = freqs_indices[k];
    a.w[k] = p->rate_weights[k];
  }
  const unsigned int spb = AN_THREADS / a.group;
  unsigned long long blocks = ((unsigned long long)a.sites + spb - 1) / spb;
  const unsigned long long cap = (unsigned long long)c->sm_count * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  ancestral_kernel<<<(unsigned)blocks, AN_THREADS, 0, c->stream>>>(a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "ancestral kernel launch");
}
