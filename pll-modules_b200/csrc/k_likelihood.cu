/*
 * k_likelihood.cu — edge and root log-likelihood with on-device reduction.
 *
 * Replaces pll_core_edge_loglikelihood_ii / _ti{,_4x4} and pll_core_root_loglikelihood
 * (libs/libpll/src/core_likelihood.c:25-190, 352-610, 1192-1497 and SIMD twins), driven like
 * likelihood.c:122-189 and :586-636.
 *
 * Per site n:
 *   terma   = sum_k w_k (1-pinv_k) sum_i P[n,k,i] pi_k[i] sum_j M_k[i][j] C[n,k,j]
 *   terminv = sum_k w_k pinv_k pi_k[inv[n]]            (inv[n] != -1, pinv_k > 0)
 *   no scaling      : ln(terma + terminv)
 *   c scalings > 0  : terminv > 0 ? ln(terma * 2^(-256 min(c,4)) + terminv) : ln(terma) + c ln 2^-256
 *   per-rate scalers: c_k normalised to their minimum, differences capped at 4 (core_likelihood.c:1396-1414)
 * times pattern_weights[n]; optional per-site output; deterministic grid sum (reduce.cuh).
 *
 * Thread = (site, rate category); the K lanes of a site combine with shuffles (4-state path) or
 * shared memory (generic path).  HBM-bound: 268 B/site (ii), 137 B/site (ti) for DNA Gamma4.
 */
#include "internal.h"
#include "devutil.cuh"
#include "reduce.cuh"

#include <utility>

#define LK_MAXK 32

struct LoglArgs
{
  const double * pclv;          /* inner ("parent") CLV */
  const double * cclv;          /* child CLV, null if tip pattern / root */
  const unsigned char * ctip;   /* child tip codes */
  const double * pmat;          /* [K][s][sp]; null for root */
  const unsigned int * pscaler;
  const unsigned int * cscaler;
  const unsigned int * weights;
  const int * invariant;        /* may be null */
  const double * eigen;         /* model block (frequencies live at +2*s*sp+sp) */
  const unsigned long long * tipmap;
  double * persite;             /* may be null */
  unsigned long long eigen_stride;
  unsigned int sites, states, sp, K;
  int rate_scalers;
  int is4;
  unsigned int freq_idx[LK_MAXK];
  double w[LK_MAXK];
  double pinv[LK_MAXK];
  ReduceOut red;
};

__device__ __forceinline__ double site_finish(double terma, double terminv, unsigned int scalings,
                                              bool root)
{
  /* root: ln(term) + c ln 2^-256 irrespective of invariant sites (core_likelihood.c:172-176) */
  if (root) return log(terma + terminv) + (double)scalings * LN_SCALE_THRESHOLD;
  if (scalings)
  {
    if (terminv > 0.0)
    {
      const unsigned int capped = scalings < PLL_SCALE_RATE_MAXDIFF ? scalings : PLL_SCALE_RATE_MAXDIFF;
      double f = SCALE_THRESHOLD_D;
      for (unsigned int q = 1; q < capped; ++q) f *= SCALE_THRESHOLD_D;
      return log(terma * f + terminv);
    }
    return log(terma) + (double)scalings * LN_SCALE_THRESHOLD;
  }
  return log(terma + terminv);
}

#define LK_THREADS 256
#define LK_U 2 /* elements per thread and trip */

/* ---- 4-state path: ROOT=false edge (ii or ti), ROOT=true root ----------------------------- */
template <int K, bool ROOT>
__global__ void __launch_bounds__(LK_THREADS, 3) logl_dna_kernel(LoglArgs a)
{
  __shared__ double s_mat[16 * K];
  __shared__ double s_freq[4 * K];
  if (!ROOT)
    for (int i = threadIdx.x; i < 16 * K; i += blockDim.x) s_mat[i] = a.pmat[i];
  for (int i = threadIdx.x; i < 4 * K; i += blockDim.x)
    s_freq[i] = a.eigen[(unsigned long long)a.freq_idx[i / 4] * a.eigen_stride + 2 * 4 * 4 + 4 + (i & 3)];
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const int k = lane % K;
  const double wk = a.w[k], pinv = a.pinv[k];
  const double f0 = s_freq[k * 4 + 0], f1 = s_freq[k * 4 + 1], f2 = s_freq[k * 4 + 2],
               f3 = s_freq[k * 4 + 3];
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  const unsigned long long Eround = (E + 31) / 32 * 32;
  double acc = 0.0;

  /* 268 B (ii) / 137 B (ti) per site: a stream.  LK_U elements per thread and trip, every load of a trip (both
   * CLV elements, scaler counts, pattern weight, tip code) issued before the first use — one element per trip
   * left the kernel latency-bound at 2.1 TB/s (round 1); the grid is one resident wave. */
  const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long trips = (Eround + stride - 1) / stride;
  for (unsigned long long t0 = 0; t0 < trips; t0 += LK_U)
  {
    d4 pv[LK_U], cv[LK_U];
    unsigned int scv[LK_U], wv[LK_U], codev[LK_U];
    unsigned long long ee[LK_U];
#pragma unroll
    for (int u = 0; u < LK_U; ++u)
    {
      ee[u] = tid + (t0 + u < trips ? t0 + u : trips) * stride;
      pv[u] = cv[u] = d4{0.0, 0.0, 0.0, 0.0};
      scv[u] = wv[u] = codev[u] = 0;
      if (ee[u] < E)
      {
        const unsigned long long site = ee[u] / K;
        pv[u] = ldg256(a.pclv + ee[u] * 4);
        if (!ROOT)
        {
          if (a.ctip)
            codev[u] = a.ctip[site];
          else
            cv[u] = ldg256(a.cclv + ee[u] * 4);
        }
        if (a.rate_scalers)
        {
          if (a.pscaler) scv[u] += a.pscaler[ee[u]];
          if (a.cscaler) scv[u] += a.cscaler[ee[u]];
        }
        else if (k == 0)
        {
          if (a.pscaler) scv[u] += a.pscaler[site];
          if (a.cscaler) scv[u] += a.cscaler[site];
        }
        if (k == 0) wv[u] = __ldg(a.weights + site);
      }
    }
#pragma unroll
    for (int u = 0; u < LK_U; ++u)
    {
      if (ee[u] >= Eround) continue; /* warp-uniform */
      const bool act = ee[u] < E;
      const unsigned long long site = ee[u] / K;
      double term = 0.0, tinv = 0.0;
      const unsigned int sc = scv[u];
      if (act)
      {
        const d4 p = pv[u];
        double t0_, t1, t2, t3;
        if (ROOT)
        {
          t0_ = t1 = t2 = t3 = 1.0;
        }
        else if (a.ctip)
        {
          const unsigned int code = codev[u];
          const double * m = s_mat + k * 16;
          t0_ = ((code & 1) ? m[0] : 0.0) + ((code & 2) ? m[1] : 0.0) + ((code & 4) ? m[2] : 0.0) + ((code & 8) ? m[3] : 0.0);
          t1 = ((code & 1) ? m[4] : 0.0) + ((code & 2) ? m[5] : 0.0) + ((code & 4) ? m[6] : 0.0) + ((code & 8) ? m[7] : 0.0);
          t2 = ((code & 1) ? m[8] : 0.0) + ((code & 2) ? m[9] : 0.0) + ((code & 4) ? m[10] : 0.0) + ((code & 8) ? m[11] : 0.0);
          t3 = ((code & 1) ? m[12] : 0.0) + ((code & 2) ? m[13] : 0.0) + ((code & 4) ? m[14] : 0.0) + ((code & 8) ? m[15] : 0.0);
        }
        else
        {
          const d4 c = cv[u];
          const double * m = s_mat + k * 16;
          t0_ = m[0] * c.x + m[1] * c.y + m[2] * c.z + m[3] * c.w;
          t1 = m[4] * c.x + m[5] * c.y + m[6] * c.z + m[7] * c.w;
          t2 = m[8] * c.x + m[9] * c.y + m[10] * c.z + m[11] * c.w;
          t3 = m[12] * c.x + m[13] * c.y + m[14] * c.z + m[15] * c.w;
        }
        term = p.x * f0 * t0_ + p.y * f1 * t1 + p.z * f2 * t2 + p.w * f3 * t3;
      }
      unsigned int site_sc = sc;
      if (a.rate_scalers)
      {
        unsigned int mn = sc;
#pragma unroll
        for (int off = 1; off < K; off <<= 1)
        {
          const unsigned int o = __shfl_xor_sync(0xffffffffu, mn, off);
          mn = o < mn ? o : mn;
        }
        unsigned int diff = sc - mn;
        if (diff > PLL_SCALE_RATE_MAXDIFF) diff = PLL_SCALE_RATE_MAXDIFF;
        for (unsigned int q = 0; q < diff; ++q) term *= SCALE_THRESHOLD_D;
        site_sc = mn;
      }
      if (pinv > 0.0)
      {
        term = wk * term * (1.0 - pinv);
        if (act && a.invariant)
        {
          const int inv = a.invariant[site];
          if (inv >= 0) tinv = wk * s_freq[k * 4 + inv] * pinv;
        }
      }
      else
        term = term * wk;
#pragma unroll
      for (int off = 1; off < K; off <<= 1)
      {
        term += __shfl_xor_sync(0xffffffffu, term, off);
        tinv += __shfl_xor_sync(0xffffffffu, tinv, off);
      }
      if (act && k == 0)
      {
        double lk = site_finish(term, tinv, site_sc, ROOT);
        lk *= (double)wv[u];
        if (a.persite) a.persite[site] = lk;
        acc += lk;
      }
    }
  }
  double v[1] = {acc};
  grid_reduce_finish<1>(v, a.red);
}

/* ---- 4 states, 4 categories, per-site scalers, no invariant sites: the hot shape (BASELINE C2/C4/C5) --------
 * The kernel above spends 290 warp instructions per 32 elements, most of them on its generality and on a
 * log() that runs with a quarter of the lanes (profiles/r02_evals_ncu_summary.csv: 93 us per 1 M sites, DRAM at
 * 18 %).  Here a warp takes a tile of 32 sites = 128 elements as 4 rows of 32: a thread's 4 elements belong to 4
 * different sites and one rate.  All 8 256-bit loads of a trip are in flight before the first use; register slot j
 * holds row j ^ perm (perm = bit-reversed rate index), so that two exchange rounds with fixed send/keep slots leave
 * in slot 0 the sum over the rates of a DIFFERENT site in each lane (the trick of the Newton kernel, k_newton.cu):
 * 3 shuffles of doubles per 4 elements instead of 8, and the logarithm, scaler term, weight and per-site store run
 * on all 32 lanes.  Same arithmetic per element and same order over the rates as logl_dna_kernel<4>: per-site
 * values are bit-identical, only the order of the final sum differs. */
int g_opt_fast_evals = 1;
#define LKF_THREADS 256
template <bool CTIP>
__global__ void __launch_bounds__(LKF_THREADS, 2) logl_dna4_fast_kernel(LoglArgs a)
{
  __shared__ __align__(16) double s_mat[4 * 18];
  __shared__ __align__(16) double s_freq[16];
  for (int i = threadIdx.x; i < 64; i += blockDim.x) s_mat[(i >> 4) * 18 + (i & 15)] = a.pmat[i];
  for (int i = threadIdx.x; i < 16; i += blockDim.x)
    s_freq[i] = a.eigen[(unsigned long long)a.freq_idx[i / 4] * a.eigen_stride + 2 * 4 * 4 + 4 + (i & 3)];
  __syncthreads();
  const unsigned int lane = threadIdx.x & 31u, k = lane & 3u;
  const unsigned int perm = ((k & 1u) << 1) | (k >> 1);
  const double wk = a.w[k];
  /* matrix and frequencies are re-read from shared memory per element (rows of different rates in different
   * banks): the 128 registers hold the 8 loads of a trip */
  const double * m = s_mat + k * 18;
  const double * f = s_freq + k * 4;
  const unsigned long long tiles = ((unsigned long long)a.sites + 31) / 32;
  const unsigned long long nwarps = (unsigned long long)gridDim.x * (LKF_THREADS / 32);
  double acc = 0.0;
  for (unsigned long long tile = (unsigned long long)blockIdx.x * (LKF_THREADS / 32) + (threadIdx.x >> 5); tile < tiles;
       tile += nwarps)
  {
    const unsigned long long site0 = tile * 32 + (lane >> 2); /* + 8 * row */
    const double * pp = a.pclv + (tile * 128 + lane) * 4;
    const double * cp = CTIP ? nullptr : a.cclv + (tile * 128 + lane) * 4;
    d4 pv[4], cv[4];
    unsigned int code[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
      const unsigned int r = (unsigned int)j ^ perm;
      pv[j] = cv[j] = d4{0.0, 0.0, 0.0, 0.0};
      code[j] = 0;
      if (site0 + 8 * r < a.sites)
      {
        pv[j] = ldg256(pp + r * 128);
        if (CTIP)
          code[j] = a.ctip[site0 + 8 * r];
        else
          cv[j] = ldg256(cp + r * 128);
      }
    }
    /* the site this lane finishes: row perm */
    const unsigned long long sitef = site0 + 8 * perm;
    const bool actf = sitef < a.sites;
    unsigned int sc = 0, wgt = 0;
    if (actf)
    {
      wgt = __ldg(a.weights + sitef);
      if (a.pscaler) sc += a.pscaler[sitef];
      if (a.cscaler) sc += a.cscaler[sitef];
    }
    double term[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
      double t0, t1, t2, t3;
      if (CTIP)
      {
        const unsigned int cd = code[j];
        t0 = ((cd & 1) ? m[0] : 0.0) + ((cd & 2) ? m[1] : 0.0) + ((cd & 4) ? m[2] : 0.0) + ((cd & 8) ? m[3] : 0.0);
        t1 = ((cd & 1) ? m[4] : 0.0) + ((cd & 2) ? m[5] : 0.0) + ((cd & 4) ? m[6] : 0.0) + ((cd & 8) ? m[7] : 0.0);
        t2 = ((cd & 1) ? m[8] : 0.0) + ((cd & 2) ? m[9] : 0.0) + ((cd & 4) ? m[10] : 0.0) + ((cd & 8) ? m[11] : 0.0);
        t3 = ((cd & 1) ? m[12] : 0.0) + ((cd & 2) ? m[13] : 0.0) + ((cd & 4) ? m[14] : 0.0) + ((cd & 8) ? m[15] : 0.0);
      }
      else
      {
        const d4 c = cv[j];
        t0 = m[0] * c.x + m[1] * c.y + m[2] * c.z + m[3] * c.w;
        t1 = m[4] * c.x + m[5] * c.y + m[6] * c.z + m[7] * c.w;
        t2 = m[8] * c.x + m[9] * c.y + m[10] * c.z + m[11] * c.w;
        t3 = m[12] * c.x + m[13] * c.y + m[14] * c.z + m[15] * c.w;
      }
      const d4 p = pv[j];
      term[j] = (p.x * f[0] * t0 + p.y * f[1] * t1 + p.z * f[2] * t2 + p.w * f[3] * t3) * wk;
    }
    term[0] += __shfl_xor_sync(0xffffffffu, term[2], 1);
    term[1] += __shfl_xor_sync(0xffffffffu, term[3], 1);
    term[0] += __shfl_xor_sync(0xffffffffu, term[1], 2);
    if (actf)
    {
      double lk = site_finish(term[0], 0.0, sc, false);
      lk *= (double)wgt;
      if (a.persite) a.persite[sitef] = lk;
      acc += lk;
    }
  }
  double v[1] = {acc};
  grid_reduce_finish<1>(v, a.red);
}

/* ---- the same shape with the CLVs streamed through a TMA ring ------------------------------------------------
 * The register kernel above is latency-bound (ncu, 1 M sites: 16 warps per SM issue their loads, wait, compute
 * ~440 instructions per trip and only then ask for more: 3.2 TB/s).  Here one producer thread keeps whole stages in
 * flight by cp.async.bulk — per 128 sites both CLV tiles (2 x 16 KB, or the tip codes), the scaler counts and the
 * pattern weights, completion on one mbarrier per stage — while 8 consumer warps lift a stage into registers, hand
 * it back and compute: the pattern of newton_dna4_tma_kernel (k_newton.cu).  Thread t owns elements t and 256 + t
 * of a tile = one rate of two sites; slot j holds row j ^ (k & 1), so that one fixed-slot exchange and one
 * butterfly leave the sum over the rates in the lanes with k < 2, one site each.  Per-site arithmetic and the
 * order over the rates are those of logl_dna_kernel<4>. */
#define LT_SITES 128
#define LT_TILE (LT_SITES * 4)
#define LT_CONSUMERS 256
#define LT_THREADS (LT_CONSUMERS + 32)

template <bool CTIP>
struct LoglTmaSmem
{
  static constexpr int STAGES = CTIP ? 5 : 3;
  double p[STAGES][LT_TILE * 4];
  double c[CTIP ? 1 : STAGES][CTIP ? 4 : LT_TILE * 4];
  unsigned int psc[STAGES][LT_SITES];
  unsigned int csc[STAGES][LT_SITES];
  unsigned int wts[STAGES][LT_SITES];
  unsigned char code[STAGES][LT_SITES];
  unsigned long long full[STAGES];
  unsigned long long empty[STAGES];
  double mat[4 * 18];
  double freq[16];
  double tab[CTIP ? 16 * 16 : 1]; /* tip child: [code][rate][row] = sum of the matrix columns in the code's mask */
};

__device__ __forceinline__ void lt_mbar_arrive(unsigned long long * bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

/* 32 bytes of a staged element without bank conflicts: lanes 4-7 of every 8 read the upper half first */
__device__ __forceinline__ d4 lt_lift(const double * src, unsigned int swz)
{
  const double2 h0 = *reinterpret_cast<const double2 *>(src + 2 * swz);
  const double2 h1 = *reinterpret_cast<const double2 *>(src + 2 * (swz ^ 1u));
  const double2 lo = swz ? h1 : h0, hi = swz ? h0 : h1;
  return d4{lo.x, lo.y, hi.x, hi.y};
}

template <bool CTIP>
__global__ void __launch_bounds__(LT_THREADS, 2) logl_dna4_tma_kernel(LoglArgs a)
{
  using Smem = LoglTmaSmem<CTIP>;
  constexpr int STAGES = Smem::STAGES;
  extern __shared__ __align__(128) unsigned char lt_smem_raw[];
  Smem & sm = *reinterpret_cast<Smem *>(lt_smem_raw);
  const unsigned int tid = threadIdx.x, lane = tid & 31u;
  const bool producer = tid >= LT_CONSUMERS;
  if (tid == 0)
  {
    for (int st = 0; st < STAGES; ++st)
    {
      mbar_init(&sm.full[st], 1);
      mbar_init(&sm.empty[st], LT_CONSUMERS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (unsigned int i = tid; i < 64; i += blockDim.x) sm.mat[(i >> 4) * 18 + (i & 15)] = a.pmat[i];
  for (unsigned int i = tid; i < 16; i += blockDim.x)
    sm.freq[i] = a.eigen[(unsigned long long)a.freq_idx[i / 4] * a.eigen_stride + 2 * 4 * 4 + 4 + (i & 3)];
  if (CTIP && tid < 256)
  {
    /* added in state order like the masked sums of the general kernel (absent states add 0.0: same bits) */
    const unsigned int cd = tid >> 4;
    const double * mr = a.pmat + ((tid >> 2) & 3u) * 16 + (tid & 3u) * 4;
    sm.tab[tid] = ((cd & 1) ? mr[0] : 0.0) + ((cd & 2) ? mr[1] : 0.0) + ((cd & 4) ? mr[2] : 0.0) + ((cd & 8) ? mr[3] : 0.0);
  }
  __syncthreads();

  const unsigned long long tiles = ((unsigned long long)a.sites + LT_SITES - 1) / LT_SITES;
  const unsigned int G = gridDim.x, b = blockIdx.x;
  const unsigned int nt = b < tiles ? (unsigned int)((tiles - b + G - 1) / G) : 0u;
  double acc = 0.0;

  if (producer)
  {
    if (lane == 0)
      for (unsigned int i = 0; i < nt; ++i)
      {
        const unsigned int slot = i % STAGES, fill = i / STAGES;
        if (fill) mbar_wait(&sm.empty[slot], (fill - 1u) & 1u); /* previous contents are in registers */
        const unsigned long long site0 = ((unsigned long long)b + (unsigned long long)i * G) * LT_SITES;
        const unsigned int ns = (unsigned int)((a.sites - site0) < LT_SITES ? (a.sites - site0) : LT_SITES);
        const unsigned int b4 = ((ns + 3u) & ~3u) * 4u;  /* rows are padded to 256 bytes: rounding up stays inside */
        const unsigned int b1 = (ns + 15u) & ~15u;
        unsigned int bytes = ns * 128u + b4;
        if (CTIP) bytes += b1; else bytes += ns * 128u;
        if (a.pscaler) bytes += b4;
        if (a.cscaler) bytes += b4;
        mbar_expect_tx(&sm.full[slot], bytes);
        tma_load_1d(sm.p[slot], a.pclv + site0 * 16, ns * 128u, &sm.full[slot]);
        if (CTIP)
          tma_load_1d(sm.code[slot], a.ctip + site0, b1, &sm.full[slot]);
        else
          tma_load_1d(sm.c[slot], a.cclv + site0 * 16, ns * 128u, &sm.full[slot]);
        tma_load_1d(sm.wts[slot], a.weights + site0, b4, &sm.full[slot]);
        if (a.pscaler) tma_load_1d(sm.psc[slot], a.pscaler + site0, b4, &sm.full[slot]);
        if (a.cscaler) tma_load_1d(sm.csc[slot], a.cscaler + site0, b4, &sm.full[slot]);
      }
  }
  else
  {
    const unsigned int k = tid & 3u, row0 = k & 1u;
    const unsigned int swz = (tid >> 2) & 1u;
    const double wk = a.w[k];
    const double * m = sm.mat + k * 18;
    const double * f = sm.freq + k * 4;
    const bool finisher_lane = (k & 2u) == 0u;
    for (unsigned int i = 0; i < nt; ++i)
    {
      const unsigned int slot = i % STAGES;
      const unsigned long long site0 = ((unsigned long long)b + (unsigned long long)i * G) * LT_SITES;
      const unsigned int ns = (unsigned int)((a.sites - site0) < LT_SITES ? (a.sites - site0) : LT_SITES);
      mbar_wait(&sm.full[slot], (i / STAGES) & 1u);
      d4 pv[2], cv[2];
      unsigned int code[2];
#pragma unroll
      for (int j = 0; j < 2; ++j)
      {
        const unsigned int el = (((unsigned int)j ^ row0) << 8) + tid;
        pv[j] = lt_lift(sm.p[slot] + (size_t)el * 4, swz);
        if (CTIP)
          code[j] = sm.code[slot][el >> 2];
        else
          cv[j] = lt_lift(sm.c[slot] + (size_t)el * 4, swz);
      }
      const unsigned int sl = (row0 << 6) + (tid >> 2); /* tile-local site this lane finishes */
      const bool actf = finisher_lane && sl < ns;
      unsigned int wgt = sm.wts[slot][sl], sc = 0;
      if (a.pscaler) sc += sm.psc[slot][sl];
      if (a.cscaler) sc += sm.csc[slot][sl];
      __syncwarp();
      if (lane == 0) lt_mbar_arrive(&sm.empty[slot]); /* this warp no longer needs the stage */
      double term[2];
#pragma unroll
      for (int j = 0; j < 2; ++j)
      {
        const unsigned int el = (((unsigned int)j ^ row0) << 8) + tid;
        double t0, t1, t2, t3;
        if (CTIP)
        {
          const double * tr = sm.tab + ((code[j] & 15u) << 4) + (k << 2);
          const double2 ta = *reinterpret_cast<const double2 *>(tr), tb = *reinterpret_cast<const double2 *>(tr + 2);
          t0 = ta.x;
          t1 = ta.y;
          t2 = tb.x;
          t3 = tb.y;
        }
        else
        {
          const d4 c = cv[j];
          t0 = m[0] * c.x + m[1] * c.y + m[2] * c.z + m[3] * c.w;
          t1 = m[4] * c.x + m[5] * c.y + m[6] * c.z + m[7] * c.w;
          t2 = m[8] * c.x + m[9] * c.y + m[10] * c.z + m[11] * c.w;
          t3 = m[12] * c.x + m[13] * c.y + m[14] * c.z + m[15] * c.w;
        }
        const d4 p = pv[j];
        term[j] = (p.x * f[0] * t0 + p.y * f[1] * t1 + p.z * f[2] * t2 + p.w * f[3] * t3) * wk;
        if ((el >> 2) >= ns) term[j] = 0.0; /* last, partial tile: what lies behind it in the stage is stale */
      }
      term[0] += __shfl_xor_sync(0xffffffffu, term[1], 1);
      term[0] += __shfl_xor_sync(0xffffffffu, term[0], 2);
      if (actf)
      {
        double lk = site_finish(term[0], 0.0, sc, false);
        lk *= (double)wgt;
        if (a.persite) a.persite[site0 + sl] = lk;
        acc += lk;
      }
    }
  }
  double v[1] = {acc};
  grid_reduce_finish<1>(v, a.red);
}

/* ---- 20-state path (protein) ------------------------------------------------------------------
 * thread = (site, category) like the 4-state kernel; the child's 20 values (or the tip mask as 0/1) and
 * the parent's sit in registers (5 + 5 256-bit loads), the K P-matrices in shared memory with a row
 * pitch that puts the K categories on different banks, so that the 10 LDS.128 of a matrix row are
 * conflict-free broadcasts.  400 FMA + 200 LDS.128 per (site, category): 1292 B/site (ii) stay the
 * bound at ~6 TB/s (the generic kernel re-read matrix and child from L1 400 times: 0.3 TB/s). */
#define LKA_THREADS 128
#define LKA_KPITCH 404 /* doubles between the matrices of two categories: 808 words = 8 banks apart */

template <int K, bool ROOT>
__global__ void __launch_bounds__(LKA_THREADS) logl_aa_kernel(LoglArgs a)
{
  __shared__ __align__(16) double s_mat[K * LKA_KPITCH];
  __shared__ double s_freq[K * 20];
  if (!ROOT)
    for (int i = threadIdx.x; i < 400 * K; i += blockDim.x) s_mat[(i / 400) * LKA_KPITCH + i % 400] = a.pmat[i];
  for (int i = threadIdx.x; i < 20 * K; i += blockDim.x)
    s_freq[i] = a.eigen[(unsigned long long)a.freq_idx[i / 20] * a.eigen_stride + 2 * 20 * 20 + 20 + i % 20];
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const int k = lane % K;
  const double wk = a.w[k], pinv = a.pinv[k];
  const double * fk = s_freq + k * 20;
  const double * mk = s_mat + k * LKA_KPITCH;
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  const unsigned long long Eround = (E + 31) / 32 * 32;
  double acc = 0.0;

  for (unsigned long long e = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; e < Eround;
       e += stride)
  {
    const bool act = e < E;
    const unsigned long long site = e / K;
    double term = 0.0, tinv = 0.0;
    unsigned int sc = 0;
    if (act)
    {
      d4 p[5], c[5];
#pragma unroll
      for (int q = 0; q < 5; ++q) p[q] = ldg256(a.pclv + e * 20 + q * 4);
      if (!ROOT)
      {
        if (a.ctip)
        {
          const unsigned long long m = a.tipmap[a.ctip[site]];
#pragma unroll
          for (int q = 0; q < 5; ++q)
          {
            c[q].x = (double)((m >> (4 * q)) & 1ull);
            c[q].y = (double)((m >> (4 * q + 1)) & 1ull);
            c[q].z = (double)((m >> (4 * q + 2)) & 1ull);
            c[q].w = (double)((m >> (4 * q + 3)) & 1ull);
          }
        }
        else
        {
#pragma unroll
          for (int q = 0; q < 5; ++q) c[q] = ldg256(a.cclv + e * 20 + q * 4);
        }
      }
      const double * pp = reinterpret_cast<const double *>(p);
#pragma unroll
      for (int i = 0; i < 20; ++i)
      {
        double tb = 1.0;
        if (!ROOT)
        {
          const d2 * row = reinterpret_cast<const d2 *>(mk + i * 20);
          double t0 = 0.0, t1 = 0.0;
#pragma unroll
          for (int q = 0; q < 5; ++q)
          {
            const d2 r0 = row[2 * q], r1 = row[2 * q + 1];
            t0 += r0.x * c[q].x + r0.y * c[q].y;
            t1 += r1.x * c[q].z + r1.y * c[q].w;
          }
          tb = t0 + t1;
        }
        term += pp[i] * fk[i] * tb;
      }
      if (k == 0)
      {
        if (a.pscaler) sc += a.pscaler[site];
        if (a.cscaler) sc += a.cscaler[site];
      }
    }
    if (pinv > 0.0)
    {
      term = wk * term * (1.0 - pinv);
      if (act && a.invariant)
      {
        const int inv = a.invariant[site];
        if (inv >= 0) tinv = wk * fk[inv] * pinv;
      }
    }
    else
      term = term * wk;
#pragma unroll
    for (int off = 1; off < K; off <<= 1)
    {
      term += __shfl_xor_sync(0xffffffffu, term, off);
      tinv += __shfl_xor_sync(0xffffffffu, tinv, off);
    }
    if (act && k == 0)
    {
      double lk = site_finish(term, tinv, sc, ROOT);
      lk *= (double)a.weights[site];
      if (a.persite) a.persite[site] = lk;
      acc += lk;
    }
  }
  double v[1] = {acc};
  grid_reduce_finish<1>(v, a.red);
}

/* ---- generic path --------------------------------------------------------------------------- */
#define LKG_THREADS 128

__global__ void __launch_bounds__(LKG_THREADS) logl_generic_kernel(LoglArgs a, int root)
{
  __shared__ double s_term[LKG_THREADS];
  __shared__ double s_inv[LKG_THREADS];
  __shared__ unsigned int s_sc[LKG_THREADS];
  const unsigned int K = a.K, s = a.states, sp = a.sp;
  const unsigned int spb = LKG_THREADS / K;
  const unsigned int lsite = threadIdx.x / K, k = threadIdx.x % K;
  const unsigned long long nblk = ((unsigned long long)a.sites + spb - 1) / spb;
  double acc = 0.0;

  for (unsigned long long blk = blockIdx.x; blk < nblk; blk += gridDim.x)
  {
    const unsigned long long site = blk * spb + lsite;
    const bool act = lsite < spb && site < a.sites;
    double term = 0.0, tinv = 0.0;
    unsigned int sc = 0;
    if (act)
    {
      const unsigned long long e = site * K + k;
      const double * freqs = a.eigen + (unsigned long long)a.freq_idx[k] * a.eigen_stride + 2 * s * sp + sp;
      const double * pv = a.pclv + e * sp;
      const double * mat = root ? nullptr : a.pmat + (size_t)k * s * sp;
      const double * cv = a.cclv ? a.cclv + e * sp : nullptr;
      unsigned long long cmask = 0;
      if (a.ctip)
      {
        const unsigned int code = a.ctip[site];
        cmask = a.is4 ? (unsigned long long)code : a.tipmap[code];
      }
      for (unsigned int i = 0; i < s; ++i)
      {
        double tb = 1.0;
        if (!root)
        {
          tb = 0.0;
          if (cv)
            for (unsigned int j = 0; j < s; ++j) tb += __ldg(mat + i * sp + j) * cv[j];
          else
            for (unsigned int j = 0; j < s; ++j)
              if ((cmask >> j) & 1ull) tb += __ldg(mat + i * sp + j);
        }
        term += pv[i] * freqs[i] * tb;
      }
      const unsigned long long si = a.rate_scalers ? e : site;
      if (a.pscaler) sc += a.pscaler[si];
      if (a.cscaler) sc += a.cscaler[si];
      const double pinv = a.pinv[k];
      if (pinv > 0.0)
      {
        const int inv = a.invariant ? a.invariant[site] : -1;
        if (inv >= 0) tinv = a.w[k] * freqs[inv] * pinv;
      }
    }
    s_term[threadIdx.x] = term;
    s_inv[threadIdx.x] = tinv;
    s_sc[threadIdx.x] = sc;
    __syncthreads();
    if (act && k == 0)
    {
      unsigned int site_sc = s_sc[threadIdx.x];
      if (a.rate_scalers)
        for (unsigned int q = 1; q < K; ++q)
          if (s_sc[threadIdx.x + q] < site_sc) site_sc = s_sc[threadIdx.x + q];
      double ta = 0.0, ti = 0.0;
      for (unsigned int q = 0; q < K; ++q)
      {
        double t = s_term[threadIdx.x + q];
        if (a.rate_scalers)
        {
          unsigned int diff = s_sc[threadIdx.x + q] - site_sc;
          if (diff > PLL_SCALE_RATE_MAXDIFF) diff = PLL_SCALE_RATE_MAXDIFF;
          for (unsigned int z = 0; z < diff; ++z) t *= SCALE_THRESHOLD_D;
        }
        const double pinv = a.pinv[q];
        if (pinv > 0.0)
          ta += a.w[q] * t * (1.0 - pinv);
        else
          ta += t * a.w[q];
        ti += s_inv[threadIdx.x + q];
      }
      double lk = site_finish(ta, ti, site_sc, root != 0);
      lk *= (double)a.weights[site];
      if (a.persite) a.persite[site] = lk;
      acc += lk;
    }
    __syncthreads();
  }
  double v[1] = {acc};
  grid_reduce_finish<1>(v, a.red);
}

/* ---- ascertainment-bias correction ------------------------------------------------------------
 * likelihood.c:24-120 (root), :191-268 (tip-inner), :342-440 (inner-inner): the partition carries
 * `states` extra "sites" after the real ones (site sites+n = every tip in state n).  Their
 * likelihoods l_n (unscaled: times 2^(-256 c_n)) correct the sum of the main kernel:
 *   Lewis        logL -= W ln(1 - sum_n l_n)           W = pattern_weight_sum
 *   Felsenstein  logL += (sum_n w_n) ln(sum_n l_n)     w_n = pattern_weights[sites+n]
 *   Stamatakis   logL += sum_n [ w_n ln l'_n + c_n ln 2^-256 ]   (l'_n still scaled; the reference
 *                does not weight the scaling term)
 * One CTA, thread n = state n; runs on the same stream right after the reducing kernel and rewrites
 * its result slot (device copy and mapped host copy). */
#define ASC_THREADS 64

__global__ void __launch_bounds__(ASC_THREADS) asc_logl_kernel(LoglArgs a, int root, int asc_type,
                                                              unsigned int weight_sum)
{
  __shared__ double s_term[ASC_THREADS];
  __shared__ unsigned int s_sc[ASC_THREADS];
  const unsigned int K = a.K, s = a.states, sp = a.sp, n = threadIdx.x;
  if (n < s)
  {
    const unsigned long long site = (unsigned long long)a.sites + n;
    double term = 0.0;
    unsigned long long cmask = 0;
    if (a.ctip)
    {
      const unsigned int code = a.ctip[site];
      cmask = a.is4 ? (unsigned long long)code : a.tipmap[code];
    }
    for (unsigned int k = 0; k < K; ++k)
    {
      const unsigned long long e = site * K + k;
      const double * freqs = a.eigen + (unsigned long long)a.freq_idx[k] * a.eigen_stride + 2 * s * sp + sp;
      const double * pv = a.pclv + e * sp;
      const double * mat = root ? nullptr : a.pmat + (size_t)k * s * sp;
      const double * cv = a.cclv ? a.cclv + e * sp : nullptr;
      double term_r = 0.0;
      for (unsigned int i = 0; i < s; ++i)
      {
        double tb = 1.0;
        if (!root)
        {
          tb = 0.0;
          if (cv)
            for (unsigned int j = 0; j < s; ++j) tb += mat[i * sp + j] * cv[j];
          else
            for (unsigned int j = 0; j < s; ++j)
              if ((cmask >> j) & 1ull) tb += mat[i * sp + j];
        }
        term_r += pv[i] * freqs[i] * tb;
      }
      term += term_r * a.w[k];
    }
    unsigned int sc = 0;
    if (a.pscaler) sc += a.pscaler[site];
    if (a.cscaler) sc += a.cscaler[site];
    s_term[n] = term;
    s_sc[n] = sc;
  }
  __syncthreads();
  if (n != 0) return;
  double base = 0.0;
  unsigned int sum_w_inv = 0;
  for (unsigned int q = 0; q < s; ++q)
  {
    const unsigned int w = a.weights[a.sites + q];
    sum_w_inv += w;
    if (asc_type == PLL_ATTRIB_AB_STAMATAKIS)
    {
      double lk = log(s_term[q]) * (double)w;
      if (s_sc[q]) lk += (double)s_sc[q] * LN_SCALE_THRESHOLD;
      base += lk;
    }
    else
      base += scalbn(s_term[q], -256 * (int)s_sc[q]);
  }
  double corr;
  if (asc_type == PLL_ATTRIB_AB_LEWIS)
    corr = -((double)weight_sum * log(1.0 - base));
  else if (asc_type == PLL_ATTRIB_AB_FELSENSTEIN)
    corr = (double)sum_w_inv * log(base);
  else
    corr = base;
  const double v = a.red.result_dev[0] + corr;
  a.red.result_dev[0] = v;
  a.red.result_host[0] = v;
  publish_tag(a.red);
}

/* with an ascertainment-bias correction the correction kernel is the one that completes the slot */
static LoglArgs without_tag_if_asc(const PartCtx * c, LoglArgs a)
{
  if (c->pub.attributes & PLL_ATTRIB_AB_MASK) a.red.tag_host = nullptr;
  return a;
}

static bool launch_asc_logl(PartCtx * c, const LoglArgs & a, int root)
{
  const pll_partition_t * p = &c->pub;
  const int asc_type = (int)(p->attributes & PLL_ATTRIB_AB_MASK);
  if (!asc_type) return true;
  if (p->states > ASC_THREADS || (p->attributes & PLL_ATTRIB_RATE_SCALERS))
  {
    pllb200_set_error(PLL_ERROR_AB_NOSUPPORT,
                      "Ascertainment bias correction needs states <= %d and per-site scalers.", ASC_THREADS);
    return false;
  }
  asc_logl_kernel<<<1, ASC_THREADS, 0, c->stream>>>(a, root, asc_type, p->pattern_weight_sum);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "asc logl kernel launch");
}

/* ---- host ------------------------------------------------------------------------------------ */

static bool fill_common(PartCtx * c, LoglArgs & a, const unsigned int * freqs_indices,
                        double * d_persite, int slot)
{
  pll_partition_t * p = &c->pub;
  if (p->rate_cats > LK_MAXK)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "rate_cats > %d is not supported.", LK_MAXK);
    return false;
  }
  if (!dev_sync_inputs(c) || !dev_sync_eigen(c, freqs_indices)) return false;
  memset(&a, 0, sizeof(a));
  a.weights = c->d_weights;
  a.eigen = c->d_eigen;
  a.eigen_stride = c->eigen_stride;
  a.tipmap = c->d_tipmap;
  a.persite = d_persite;
  a.sites = p->sites;
  a.states = p->states;
  a.sp = p->states_padded;
  a.K = p->rate_cats;
  a.rate_scalers = (p->attributes & PLL_ATTRIB_RATE_SCALERS) != 0;
  a.is4 = p->states == 4;
  bool any_pinv = false;
  for (unsigned int k = 0; k < p->rate_cats; ++k)
  {
    a.freq_idx[k] = freqs_indices[k];
    a.w[k] = p->rate_weights[k];
    a.pinv[k] = p->prop_invar[freqs_indices[k]];
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
  dev_slot_tag(c, slot, &a.red.tag_host, &a.red.tag);
  return true;
}

template <bool ROOT>
static bool dispatch_logl(PartCtx * c, const LoglArgs & a)
{
  const unsigned int K = a.K;
  eval_timing_begin(c->stream);
  bool no_pinv = true;
  for (unsigned int k = 0; k < K && k < LK_MAXK; ++k) no_pinv &= !(a.pinv[k] > 0.0);
  if (!ROOT && g_opt_fast_evals && a.states == 4 && K == 4 && !a.rate_scalers && no_pinv && (a.ctip || a.cclv))
  {
    unsigned long long blocks = (((unsigned long long)a.sites + 31) / 32 + LKF_THREADS / 32 - 1) / (LKF_THREADS / 32);
    const unsigned long long cap = (unsigned long long)c->sm_count * 2; /* one resident wave */
    if (blocks > cap) blocks = cap;
    if (blocks > PLLB200_MAX_RED_BLOCKS) blocks = PLLB200_MAX_RED_BLOCKS;
    if (blocks < 1) blocks = 1;
    if (g_opt_fast_evals == 1 && a.sites >= 4096)
    {
      /* streamed through shared memory by TMA */
      static bool optin = false;
      if (!optin)
      {
        if (!pllb200_cuda_ok(cudaFuncSetAttribute(logl_dna4_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                  (int)sizeof(LoglTmaSmem<true>) + 128), "logl TMA kernel shared memory") ||
            !pllb200_cuda_ok(cudaFuncSetAttribute(logl_dna4_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                  (int)sizeof(LoglTmaSmem<false>) + 128), "logl TMA kernel shared memory"))
          return false;
        optin = true;
      }
      unsigned long long tb = ((unsigned long long)a.sites + LT_SITES - 1) / LT_SITES;
      if (tb > cap) tb = cap;
      if (tb > PLLB200_MAX_RED_BLOCKS) tb = PLLB200_MAX_RED_BLOCKS;
      if (a.ctip)
        logl_dna4_tma_kernel<true><<<(unsigned)tb, LT_THREADS, sizeof(LoglTmaSmem<true>) + 128, c->stream>>>(a);
      else
        logl_dna4_tma_kernel<false><<<(unsigned)tb, LT_THREADS, sizeof(LoglTmaSmem<false>) + 128, c->stream>>>(a);
    }
    else if (a.ctip)
      logl_dna4_fast_kernel<true><<<(unsigned)blocks, LKF_THREADS, 0, c->stream>>>(a);
    else
      logl_dna4_fast_kernel<false><<<(unsigned)blocks, LKF_THREADS, 0, c->stream>>>(a);
  }
  else if (a.states == 4 && (K == 1 || K == 2 || K == 4 || K == 8))
  {
    const unsigned long long E = (unsigned long long)a.sites * K;
    unsigned long long blocks = (E + LK_THREADS - 1) / LK_THREADS;
    const unsigned long long cap = (unsigned long long)c->sm_count * 3; /* one resident wave at 3 CTAs per SM */
    if (blocks > cap) blocks = cap;
    if (blocks > PLLB200_MAX_RED_BLOCKS) blocks = PLLB200_MAX_RED_BLOCKS;
    if (blocks < 1) blocks = 1;
    switch (K)
    {
      case 1: logl_dna_kernel<1, ROOT><<<(unsigned)blocks, LK_THREADS, 0, c->stream>>>(a); break;
      case 2: logl_dna_kernel<2, ROOT><<<(unsigned)blocks, LK_THREADS, 0, c->stream>>>(a); break;
      case 4: logl_dna_kernel<4, ROOT><<<(unsigned)blocks, LK_THREADS, 0, c->stream>>>(a); break;
      default: logl_dna_kernel<8, ROOT><<<(unsigned)blocks, LK_THREADS, 0, c->stream>>>(a); break;
    }
  }
  else if (a.states == 20 && a.sp == 20 && !a.rate_scalers && (K == 1 || K == 2 || K == 4 || K == 8))
  {
    const unsigned long long E = (unsigned long long)a.sites * K;
    unsigned long long blocks = (E + LKA_THREADS - 1) / LKA_THREADS;
    const unsigned long long cap = (unsigned long long)c->sm_count * 4;
    if (blocks > cap) blocks = cap;
    if (blocks > PLLB200_MAX_RED_BLOCKS) blocks = PLLB200_MAX_RED_BLOCKS;
    if (blocks < 1) blocks = 1;
    switch (K)
    {
      case 1: logl_aa_kernel<1, ROOT><<<(unsigned)blocks, LKA_THREADS, 0, c->stream>>>(a); break;
      case 2: logl_aa_kernel<2, ROOT><<<(unsigned)blocks, LKA_THREADS, 0, c->stream>>>(a); break;
      case 4: logl_aa_kernel<4, ROOT><<<(unsigned)blocks, LKA_THREADS, 0, c->stream>>>(a); break;
      default: logl_aa_kernel<8, ROOT><<<(unsigned)blocks, LKA_THREADS, 0, c->stream>>>(a); break;
    }
  }
  else
  {
    if (K > LKG_THREADS)
    {
      pllb200_set_error(PLL_ERROR_PARAM_INVALID, "rate_cats too large.");
      return false;
    }
    const unsigned int spb = LKG_THREADS / K;
    unsigned long long blocks = ((unsigned long long)a.sites + spb - 1) / spb;
    const unsigned long long cap = (unsigned long long)c->sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks > PLLB200_MAX_RED_BLOCKS) blocks = PLLB200_MAX_RED_BLOCKS;
    logl_generic_kernel<<<(unsigned)blocks, LKG_THREADS, 0, c->stream>>>(a, ROOT ? 1 : 0);
  }
  eval_timing_end(c->stream);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "logl kernel launch");
}

bool launch_edge_logl(PartCtx * c, unsigned int parent_clv, int parent_scaler,
                      unsigned int child_clv, int child_scaler, unsigned int matrix_index,
                      const unsigned int * freqs_indices, double * d_persite, int slot)
{
  pll_partition_t * p = &c->pub;
  if (parent_clv >= p->nodes || child_clv >= p->nodes || matrix_index >= p->prob_matrices ||
      parent_scaler >= (int)p->scale_buffers || child_scaler >= (int)p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid index in edge log-likelihood call.");
    return false;
  }
  /* the inner node plays "parent"; the matrix is applied to the other end (likelihood.c:612-624) */
  const bool ptip = is_tip_pattern(c, parent_clv), ctip = is_tip_pattern(c, child_clv);
  if (ptip && ctip)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Edge log-likelihood between two pattern tips.");
    return false;
  }
  if (ptip)
  {
    std::swap(parent_clv, child_clv);
    std::swap(parent_scaler, child_scaler);
  }
  LoglArgs a;
  if (!fill_common(c, a, freqs_indices, d_persite, slot)) return false;
  a.pclv = clv_dev(c, parent_clv);
  a.pscaler = scaler_dev(c, parent_scaler);
  if (ptip || ctip)
    a.ctip = c->d_tipchars + (size_t)child_clv * c->tip_stride;
  else
  {
    a.cclv = clv_dev(c, child_clv);
    a.cscaler = scaler_dev(c, child_scaler);
  }
  a.pmat = c->d_pmat + (size_t)matrix_index * c->pmat_elems;
  return dispatch_logl<false>(c, without_tag_if_asc(c, a)) && launch_asc_logl(c, a, 0);
}

bool launch_root_logl(PartCtx * c, unsigned int clv_index, int scaler_index,
                      const unsigned int * freqs_indices, double * d_persite, int slot)
{
  pll_partition_t * p = &c->pub;
  if (clv_index >= p->nodes || clv_index < c->clv_first || scaler_index >= (int)p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid index in root log-likelihood call.");
    return false;
  }
  LoglArgs a;
  if (!fill_common(c, a, freqs_indices, d_persite, slot)) return false;
  a.pclv = clv_dev(c, clv_index);
  a.pscaler = scaler_dev(c, scaler_index);
  return dispatch_logl<true>(c, without_tag_if_asc(c, a)) && launch_asc_logl(c, a, 1);
}

/* ---- edge log-likelihood without any allocation ----------
 * Callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it) get
 * pllb200_edge_loglikelihood_noalloc: the edge logL kernels above with their scratch — one partial sum per CTA,
 * the ticket of the last-CTA reduction, and the per-site log-likelihoods — in a file-scope device array instead
 * of the partition's cudaMalloc'ed buffers (d_partials / d_ticket, and d_persite which pll_compute_edge_
 * loglikelihood allocates on first use).  Every GPU that loads the library holds the array (2.1 MB) until exit.
 * All calls on a device share it and it keeps the last call's contents: the ticket is cleared on the stream
 * before every launch, partials and per-site values are fully overwritten before they are read.  Two calls
 * whose kernels overlap on the GPU (different partitions on different streams, several host threads) RACE on
 * it: serialise them. */
#define NOALLOC_MAX_SITES 262144u /* stated maximum: partitions with more sites are rejected */
#define NOALLOC_PERSITE_BYTES ((size_t)NOALLOC_MAX_SITES * sizeof(double))
#define NOALLOC_PARTIALS_BYTES ((size_t)PLLB200_MAX_RED_BLOCKS * 4 * sizeof(double))
#define NOALLOC_TICKET_BYTES 256
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_PERSITE_BYTES + NOALLOC_PARTIALS_BYTES + NOALLOC_TICKET_BYTES];

/* In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is
 * asked for once per device, with that device current, and kept here. */
#define NOALLOC_MAX_DEVICES 64
static unsigned char * g_noalloc_ptr[NOALLOC_MAX_DEVICES];

static unsigned char * noalloc_scratch_on(int device)
{
  if (device < 0 || device >= NOALLOC_MAX_DEVICES)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "pllb200_edge_loglikelihood_noalloc: device %d out of range.", device);
    return nullptr;
  }
  if (!g_noalloc_ptr[device])
  {
    void * ptr = nullptr;
    if (!pllb200_cuda_ok(cudaSetDevice(device), "cudaSetDevice") ||
        !pllb200_cuda_ok(cudaGetSymbolAddress(&ptr, g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)"))
      return nullptr;
    g_noalloc_ptr[device] = static_cast<unsigned char *>(ptr);
  }
  return g_noalloc_ptr[device];
}

extern "C" int pllb200_edge_loglikelihood_noalloc(pll_partition_t * partition, unsigned int parent_clv_index,
                                                  int parent_scaler_index, unsigned int child_clv_index,
                                                  int child_scaler_index, unsigned int matrix_index,
                                                  const unsigned int * freqs_indices, double * loglh_out,
                                                  double * persite_lnl)
{
  if (!partition || ctx_of(partition)->magic != PLLB200_MAGIC || !freqs_indices || !loglh_out)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "pllb200_edge_loglikelihood_noalloc: not a libpll_b200 partition, or a null argument.");
    return PLL_FAILURE;
  }
  PartCtx * c = ctx_of(partition);
  pll_partition_t * p = &c->pub;
  if (p->sites > NOALLOC_MAX_SITES)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID,
                      "pllb200_edge_loglikelihood_noalloc: %u sites exceed the static scratch (maximum %u sites); "
                      "use pll_compute_edge_loglikelihood.", p->sites, NOALLOC_MAX_SITES);
    return PLL_FAILURE;
  }
  if (p->rate_cats > LK_MAXK)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "rate_cats > %d is not supported.", LK_MAXK);
    return PLL_FAILURE;
  }
  /* everything the kernel reads must already be mirrored on the device: mirroring stages and may allocate */
  bool current = !c->any_tip_dirty && !c->weights_dirty && !c->tipmap_dirty && !(c->invariant_dirty && p->invariant);
  for (unsigned int k = 0; k < p->rate_cats && current; ++k)
  {
    const unsigned int m = freqs_indices[k];
    if (m >= p->rate_matrices || !p->eigen_decomp_valid[m] || c->eigen_version_dev[m] != c->eigen_version[m]) current = false;
    else if (p->prop_invar[m] > 0.0 && !c->d_invariant) current = false;
  }
  if (!current)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID,
                      "pllb200_edge_loglikelihood_noalloc: host-side inputs or model are newer than their device copies; "
                      "run an ordinary evaluation (pll_update_prob_matrices / pll_compute_edge_loglikelihood) once first.");
    return PLL_FAILURE;
  }
  unsigned char * scratch = noalloc_scratch_on(c->device);
  if (!scratch) return PLL_FAILURE;
  if (!pllb200_cuda_ok(cudaSetDevice(c->device), "cudaSetDevice")) return PLL_FAILURE;
  double * d_persite = reinterpret_cast<double *>(scratch);
  double * d_partials = reinterpret_cast<double *>(scratch + NOALLOC_PERSITE_BYTES);
  unsigned int * d_ticket = reinterpret_cast<unsigned int *>(scratch + NOALLOC_PERSITE_BYTES + NOALLOC_PARTIALS_BYTES);

  if (parent_clv_index >= p->nodes || child_clv_index >= p->nodes || matrix_index >= p->prob_matrices ||
      parent_scaler_index >= (int)p->scale_buffers || child_scaler_index >= (int)p->scale_buffers)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid index in edge log-likelihood call.");
    return PLL_FAILURE;
  }
  const bool ptip = is_tip_pattern(c, parent_clv_index), ctip = is_tip_pattern(c, child_clv_index);
  if (ptip && ctip)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Edge log-likelihood between two pattern tips.");
    return PLL_FAILURE;
  }
  if (ptip)
  {
    std::swap(parent_clv_index, child_clv_index);
    std::swap(parent_scaler_index, child_scaler_index);
  }
  const int slot = dev_alloc_slot(c);
  LoglArgs a;
  if (!fill_common(c, a, freqs_indices, persite_lnl ? d_persite : nullptr, slot)) return PLL_FAILURE;
  a.red.partials = d_partials; /* the static scratch instead of the partition's buffers */
  a.red.ticket = d_ticket;
  a.pclv = clv_dev(c, parent_clv_index);
  a.pscaler = scaler_dev(c, parent_scaler_index);
  if (ptip || ctip)
    a.ctip = c->d_tipchars + (size_t)child_clv_index * c->tip_stride;
  else
  {
    a.cclv = clv_dev(c, child_clv_index);
    a.cscaler = scaler_dev(c, child_scaler_index);
  }
  a.pmat = c->d_pmat + (size_t)matrix_index * c->pmat_elems;
  /* the ticket must start at zero whatever an earlier (possibly failed) call left behind */
  if (!pllb200_cuda_ok(cudaMemsetAsync(d_ticket, 0, NOALLOC_TICKET_BYTES, c->stream), "scratch ticket clear")) return PLL_FAILURE;
  if (!dispatch_logl<false>(c, without_tag_if_asc(c, a)) || !launch_asc_logl(c, a, 0)) return PLL_FAILURE;
  if (persite_lnl)
  {
    if (!pllb200_cuda_ok(cudaMemcpyAsync(persite_lnl, d_persite, sizeof(double) * p->sites, cudaMemcpyDeviceToHost, c->stream),
                         "persite copy"))
      return PLL_FAILURE;
    g_d2h_bytes += sizeof(double) * p->sites;
    if (!dev_flush(c)) return PLL_FAILURE;
  }
  else if (!(dev_wait_results(c) && dev_wait_slot(c, slot)))
    return PLL_FAILURE;
  g_d2h_bytes += sizeof(double);
  *loglh_out = c->h_result[slot * 4];
  return PLL_SUCCESS;
}
