/*
 * k_partials.cu — CLV ("partials") updates for an operation list.
 *
 * Replaces pll_update_partials + pll_core_update_partial_{ii,ti,tt} + pll_core_create_lookup
 * (libs/libpll/src/partials.c:24-291, core_partials.c:24-1211 and the SSE/AVX/AVX2 twins).
 *
 * B200 design — the "site-column" traversal kernel (4-state path):
 *   Every likelihood kernel is independent per alignment site, so instead of one launch per
 *   operation (the reference's loop, partials.c:245-291) a warp owns a chunk of sites and walks the
 *   WHOLE operation list for that chunk: thread = (site, rate category), one 256-bit load per child
 *   CLV element, one 256-bit store per parent element, warp-level requests are 1 KB contiguous.
 *   No thread ever reads an element another thread wrote, hence no barriers between operations and
 *   one launch per traversal of any length (full: n-2 ops; SPR/BLO: 1-10 ops).  A child CLV written
 *   a few operations earlier by the same warp is still in the 126 MB L2, so with the list scheduled
 *   depth-first/heavy-subtree-first (schedule_ops below) most child reads never reach HBM: DRAM
 *   traffic tends to "write each CLV once" rather than the algorithmic 396 B/site/op.
 *   The two P-matrices of an operation are staged per warp in shared memory (interleaved [row][cat]
 *   [col] image, 1 KB/op for Gamma4) and read as broadcast 128-bit loads.
 *   Tip children are handled in-register from the 1-byte tip code (masked row sums) — this is the
 *   tip-tip / tip-inner specialisation without the reference's per-op lookup tables
 *   (core_partials.c:1013-1210, core_partials_avx.c:1340-1395).
 *   Underflow scaling: per site, all K*4 entries < 2^-256 (strict, unscaled product) -> multiply by
 *   2^256 and bump the integer scaler (core_partials.c:729-763); the K lanes of a site vote with
 *   shuffles.  Tip-tip never rescales (core_partials.c:185-186); no scaler buffer -> no test.
 *
 * Algorithmic bytes per site and op (DNA, Gamma4, per-site scalers): ii 396, ti 265, tt 134.
 *
 * Generic path (any state count <= 64, any K <= 32): one launch per operation, thread = (site,cat).
 */
#include "internal.h"
#include "devutil.cuh"

#include <algorithm>

struct TravArgs
{
  const pll_operation_t * ops;
  unsigned int nops;
  double * clv;
  unsigned long long clv_stride;
  unsigned int clv_first;
  unsigned int * scaler;
  unsigned long long scaler_stride;
  const double * pmat_il;
  const unsigned char * tipchars;
  unsigned long long tip_stride;
  unsigned int tips_pattern; /* #tips when PATTERN_TIP, else 0 */
  unsigned int sites;        /* sites_alloc */
};

/* one side of the product for one element: x[i] = sum_j M_k[i][j] * v[j]  (inner child) */
template <int K>
__device__ __forceinline__ void side_inner(const double * sM, int k, const d4 & v, double x[4])
{
#pragma unroll
  for (int i = 0; i < 4; ++i)
  {
    const double2 a = *reinterpret_cast<const double2 *>(sM + (i * K + k) * 4);
    const double2 b = *reinterpret_cast<const double2 *>(sM + (i * K + k) * 4 + 2);
    x[i] = a.x * v.x + a.y * v.y + b.x * v.z + b.y * v.w;
  }
}
/* tip child: x[i] = sum over the states set in the 4-bit code of M_k[i][j] */
template <int K>
__device__ __forceinline__ void side_tip(const double * sM, int k, unsigned int code, double x[4])
{
#pragma unroll
  for (int i = 0; i < 4; ++i)
  {
    const double2 a = *reinterpret_cast<const double2 *>(sM + (i * K + k) * 4);
    const double2 b = *reinterpret_cast<const double2 *>(sM + (i * K + k) * 4 + 2);
    double t = 0.0;
    if (code & 1u) t += a.x;
    if (code & 2u) t += a.y;
    if (code & 4u) t += b.x;
    if (code & 8u) t += b.y;
    x[i] = t;
  }
}

#define TRAV_WARPS 8
#define TRAV_R 1 /* elements per lane per chunk */
#define TRAV_MIN_CTAS 4 /* 64 registers -> 32 resident warps per SM */

template <int K, bool RATE_SCALERS>
__global__ void __launch_bounds__(TRAV_WARPS * 32, TRAV_MIN_CTAS) trav_dna_kernel(TravArgs a)
{
  /* per warp: 2 stages x (L,R) x 16*K doubles */
  __shared__ __align__(16) double s_mat[TRAV_WARPS][2][2][16 * K];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int k = lane % K;
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long chunk_elems = 32ull * TRAV_R;
  const unsigned long long nchunks = (E + chunk_elems - 1) / chunk_elems;
  const unsigned long long gwarp = (unsigned long long)blockIdx.x * TRAV_WARPS + warp;
  const unsigned long long nwarps = (unsigned long long)gridDim.x * TRAV_WARPS;

  for (unsigned long long chunk = gwarp; chunk < nchunks; chunk += nwarps)
  {
    unsigned long long e[TRAV_R];
    bool act[TRAV_R];
#pragma unroll
    for (int r = 0; r < TRAV_R; ++r)
    {
      e[r] = chunk * chunk_elems + (unsigned long long)r * 32 + lane;
      act[r] = e[r] < E;
    }

    for (unsigned int o = 0; o < a.nops; ++o)
    {
      const uint4 * op4 = reinterpret_cast<const uint4 *>(a.ops + o);
      const uint4 oa = __ldg(op4);
      const uint4 ob = __ldg(op4 + 1);
      const unsigned int p_clv = oa.x;
      const int p_sc = (int)oa.y;
      const unsigned int c1_clv = oa.z, c1_mat = oa.w;
      const int c1_sc = (int)ob.x;
      const unsigned int c2_clv = ob.y, c2_mat = ob.z;
      const int c2_sc = (int)ob.w;
      const bool tip1 = c1_clv < a.tips_pattern;
      const bool tip2 = c2_clv < a.tips_pattern;

      /* ---- issue all global loads of this op ---- */
      double * sL = s_mat[warp][o & 1][0];
      double * sR = s_mat[warp][o & 1][1];
      const double2 * gL = reinterpret_cast<const double2 *>(a.pmat_il + (unsigned long long)c1_mat * 16 * K);
      const double2 * gR = reinterpret_cast<const double2 *>(a.pmat_il + (unsigned long long)c2_mat * 16 * K);
      double2 mL[(8 * K + 31) / 32], mR[(8 * K + 31) / 32];
#pragma unroll
      for (int u = 0; u < (8 * K + 31) / 32; ++u)
      {
        const int idx = u * 32 + lane;
        if (idx < 8 * K)
        {
          mL[u] = __ldg(gL + idx);
          mR[u] = __ldg(gR + idx);
        }
      }

      d4 vl[TRAV_R], vr[TRAV_R];
      unsigned int cl[TRAV_R], cr[TRAV_R];
      unsigned int sc[TRAV_R];
#pragma unroll
      for (int r = 0; r < TRAV_R; ++r)
      {
        cl[r] = cr[r] = 15u;
        sc[r] = 0;
        vl[r] = vr[r] = d4{0, 0, 0, 0};
        if (act[r])
        {
          const unsigned long long site = e[r] / K;
          if (tip1)
            cl[r] = a.tipchars[(unsigned long long)c1_clv * a.tip_stride + site];
          else
            vl[r] = ldg256(a.clv + (unsigned long long)(c1_clv - a.clv_first) * a.clv_stride + e[r] * 4);
          if (tip2)
            cr[r] = a.tipchars[(unsigned long long)c2_clv * a.tip_stride + site];
          else
            vr[r] = ldg256(a.clv + (unsigned long long)(c2_clv - a.clv_first) * a.clv_stride + e[r] * 4);
          if (p_sc >= 0)
          {
            const unsigned long long si = RATE_SCALERS ? e[r] : site;
            if (RATE_SCALERS || k == 0)
            {
              if (c1_sc >= 0) sc[r] += a.scaler[(unsigned long long)c1_sc * a.scaler_stride + si];
              if (c2_sc >= 0) sc[r] += a.scaler[(unsigned long long)c2_sc * a.scaler_stride + si];
            }
          }
        }
      }

      /* ---- stage matrices for the warp ---- */
#pragma unroll
      for (int u = 0; u < (8 * K + 31) / 32; ++u)
      {
        const int idx = u * 32 + lane;
        if (idx < 8 * K)
        {
          reinterpret_cast<double2 *>(sL)[idx] = mL[u];
          reinterpret_cast<double2 *>(sR)[idx] = mR[u];
        }
      }
      __syncwarp();

      /* ---- compute ---- */
      double * pdst = a.clv + (unsigned long long)(p_clv - a.clv_first) * a.clv_stride;
#pragma unroll
      for (int r = 0; r < TRAV_R; ++r)
      {
        double xl[4], xr[4];
        if (tip1)
          side_tip<K>(sL, k, cl[r], xl);
        else
          side_inner<K>(sL, k, vl[r], xl);
        if (tip2)
          side_tip<K>(sR, k, cr[r], xr);
        else
          side_inner<K>(sR, k, vr[r], xr);
        d4 out;
        out.x = xl[0] * xr[0];
        out.y = xl[1] * xr[1];
        out.z = xl[2] * xr[2];
        out.w = xl[3] * xr[3];

        if (p_sc >= 0)
        {
          int flag = 0;
          if (!(tip1 && tip2))
          {
            flag = (out.x < SCALE_THRESHOLD_D) & (out.y < SCALE_THRESHOLD_D) &
                   (out.z < SCALE_THRESHOLD_D) & (out.w < SCALE_THRESHOLD_D);
            if (!RATE_SCALERS)
            {
#pragma unroll
              for (int off = 1; off < K; off <<= 1) flag &= __shfl_xor_sync(0xffffffffu, flag, off);
            }
            if (flag)
            {
              out.x *= SCALE_FACTOR_D;
              out.y *= SCALE_FACTOR_D;
              out.z *= SCALE_FACTOR_D;
              out.w *= SCALE_FACTOR_D;
            }
          }
          if (act[r] && (RATE_SCALERS || k == 0))
          {
            const unsigned long long si = RATE_SCALERS ? e[r] : e[r] / K;
            a.scaler[(unsigned long long)p_sc * a.scaler_stride + si] = sc[r] + (unsigned int)flag;
          }
        }
        if (act[r]) stg256(pdst + e[r] * 4, out);
      }
    }
    __syncwarp();
  }
}

/* ------------------------------------------------------------------------------------------ */
/* generic path: any states, any K; one operation per launch                                   */
/* ------------------------------------------------------------------------------------------ */

struct GenOpArgs
{
  double * parent;
  const double * left;   /* null when tip pattern */
  const double * right;
  const unsigned char * ltip;
  const unsigned char * rtip;
  const double * lmat;   /* [K][s][sp] */
  const double * rmat;
  unsigned int * pscaler;
  const unsigned int * lscaler;
  const unsigned int * rscaler;
  const unsigned long long * tipmap;
  unsigned int sites, states, sp, K;
  int rate_scalers;
  int is4; /* 4-state partitions store the raw mask as the tip code */
};

#define GEN_THREADS 128
#define GEN_MAXS 64

__global__ void __launch_bounds__(GEN_THREADS) partial_generic_kernel(GenOpArgs a)
{
  __shared__ int s_flag[GEN_THREADS];
  const unsigned int K = a.K, s = a.states, sp = a.sp;
  const unsigned int spb = GEN_THREADS / K; /* sites per block */
  const unsigned int lsite = threadIdx.x / K;
  const unsigned int k = threadIdx.x % K;
  const unsigned long long site = (unsigned long long)blockIdx.x * spb + lsite;
  const bool act = lsite < spb && site < a.sites;

  double out[GEN_MAXS];
  int flag = 1;
  if (act)
  {
    const unsigned long long e = site * K + k;
    const double * lm = a.lmat + (size_t)k * s * sp;
    const double * rm = a.rmat + (size_t)k * s * sp;
    unsigned long long lmask = 0, rmask = 0;
    if (a.ltip)
    {
      const unsigned int code = a.ltip[site];
      lmask = a.is4 ? (unsigned long long)code : a.tipmap[code];
    }
    if (a.rtip)
    {
      const unsigned int code = a.rtip[site];
      rmask = a.is4 ? (unsigned long long)code : a.tipmap[code];
    }
    const double * lv = a.left ? a.left + e * sp : nullptr;
    const double * rv = a.right ? a.right + e * sp : nullptr;
    for (unsigned int i = 0; i < s; ++i)
    {
      double x = 0.0, y = 0.0;
      if (lv)
        for (unsigned int j = 0; j < s; ++j) x += __ldg(lm + i * sp + j) * lv[j];
      else
        for (unsigned int j = 0; j < s; ++j)
          if ((lmask >> j) & 1ull) x += __ldg(lm + i * sp + j);
      if (rv)
        for (unsigned int j = 0; j < s; ++j) y += __ldg(rm + i * sp + j) * rv[j];
      else
        for (unsigned int j = 0; j < s; ++j)
          if ((rmask >> j) & 1ull) y += __ldg(rm + i * sp + j);
      out[i] = x * y;
      flag &= (out[i] < SCALE_THRESHOLD_D);
    }
  }
  const bool tiptip = a.ltip && a.rtip;
  if (a.pscaler && !tiptip && !a.rate_scalers)
  {
    s_flag[threadIdx.x] = flag;
    __syncthreads();
    if (act)
      for (unsigned int q = 0; q < K; ++q) flag &= s_flag[lsite * K + q];
  }
  if (!act) return;
  const unsigned long long e = site * K + k;
  if (!a.pscaler || tiptip) flag = 0;
  if (a.pscaler)
  {
    const unsigned long long si = a.rate_scalers ? e : site;
    if (a.rate_scalers || k == 0)
    {
      unsigned int v = (unsigned int)flag;
      if (a.lscaler) v += a.lscaler[si];
      if (a.rscaler) v += a.rscaler[si];
      a.pscaler[si] = v;
    }
  }
  double * dst = a.parent + e * sp;
  for (unsigned int i = 0; i < s; ++i) dst[i] = flag ? out[i] * SCALE_FACTOR_D : out[i];
}

/* ------------------------------------------------------------------------------------------ */
/* host: validation, L2-friendly scheduling, dispatch                                          */
/* ------------------------------------------------------------------------------------------ */

/* Re-order a valid operation list depth-first with the heavier child subtree first, so that the
 * CLV a parent consumes was produced as recently as possible (maximises L2 hits in the site-column
 * kernel).  Only done when the list is a forest in which every buffer is written once and never
 * read before it is written; otherwise the caller's order is kept. */
static const pll_operation_t * schedule_ops(PartCtx * c, const pll_operation_t * ops,
                                            unsigned int count)
{
  if (!g_opt_trav_reorder || count < 3) return ops;
  const unsigned int nodes = c->pub.nodes;
  std::vector<int> & producer = c->idx_scratch;
  producer.assign(nodes, -1);
  for (unsigned int i = 0; i < count; ++i)
  {
    const pll_operation_t & o = ops[i];
    if (o.parent_clv_index >= nodes || producer[o.parent_clv_index] >= 0) return ops; /* rewritten */
    /* read-before-write hazard: a child that a LATER op produces */
    producer[o.parent_clv_index] = (int)i;
  }
  std::vector<int> size(count, 1), consumer(count, -1);
  for (unsigned int i = 0; i < count; ++i)
  {
    const unsigned int ch[2] = {ops[i].child1_clv_index, ops[i].child2_clv_index};
    for (int q = 0; q < 2; ++q)
      if (ch[q] < nodes && producer[ch[q]] >= 0)
      {
        const int pj = producer[ch[q]];
        if (pj > (int)i) return ops; /* consumes a value written later: order is semantic */
        if (consumer[pj] >= 0) return ops; /* shared subtree: keep order */
        consumer[pj] = (int)i;
      }
  }
  for (unsigned int i = 0; i < count; ++i) /* producers precede consumers: one forward pass */
    if (consumer[i] >= 0) size[consumer[i]] += size[i];

  c->op_scratch.clear();
  c->op_scratch.reserve(count);
  std::vector<std::pair<int, int>> stack; /* (op, state) */
  for (unsigned int root = 0; root < count; ++root)
  {
    if (consumer[root] >= 0) continue;
    stack.push_back({(int)root, 0});
    while (!stack.empty())
    {
      auto & top = stack.back();
      const int i = top.first;
      if (top.second == 1)
      {
        c->op_scratch.push_back(ops[i]);
        stack.pop_back();
        continue;
      }
      top.second = 1;
      int ch[2] = {-1, -1};
      const unsigned int cc[2] = {ops[i].child1_clv_index, ops[i].child2_clv_index};
      for (int q = 0; q < 2; ++q)
        if (cc[q] < nodes && producer[cc[q]] >= 0) ch[q] = producer[cc[q]];
      /* visit heavier first => it is pushed last */
      int first = ch[0], second = ch[1];
      if (first < 0 || (second >= 0 && size[second] > size[first])) std::swap(first, second);
      if (second >= 0) stack.push_back({second, 0});
      if (first >= 0) stack.push_back({first, 0});
    }
  }
  return c->op_scratch.size() == count ? c->op_scratch.data() : ops;
}

template <int K>
static bool launch_trav_dna(PartCtx * c, const TravArgs & a, bool rate_scalers)
{
  static int ctas_per_sm_cache[2] = {0, 0};
  const int ri = rate_scalers ? 1 : 0;
  if (!ctas_per_sm_cache[ri])
  {
    int nb = 0;
    cudaError_t e = rate_scalers
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, trav_dna_kernel<K, true>, TRAV_WARPS * 32, 0)
        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, trav_dna_kernel<K, false>, TRAV_WARPS * 32, 0);
    if (e != cudaSuccess || nb < 1) nb = 2;
    ctas_per_sm_cache[ri] = nb;
  }
  int per_sm = g_opt_trav_ctas_per_sm > 0 ? std::min(g_opt_trav_ctas_per_sm, ctas_per_sm_cache[ri])
                                          : ctas_per_sm_cache[ri];
  const unsigned long long E = (unsigned long long)a.sites * K;
  const unsigned long long nchunks = (E + 32 * TRAV_R - 1) / (32 * TRAV_R);
  unsigned long long blocks = (nchunks + TRAV_WARPS - 1) / TRAV_WARPS;
  const unsigned long long cap = (unsigned long long)c->sm_count * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (rate_scalers)
    trav_dna_kernel<K, true><<<(unsigned int)blocks, TRAV_WARPS * 32, 0, c->stream>>>(a);
  else
    trav_dna_kernel<K, false><<<(unsigned int)blocks, TRAV_WARPS * 32, 0, c->stream>>>(a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "trav_dna_kernel launch");
}

bool launch_partials(PartCtx * c, const pll_operation_t * ops_in, unsigned int count)
{
  pll_partition_t * p = &c->pub;
  if (!count) return true;
  const bool pattern_tip = (p->attributes & PLL_ATTRIB_PATTERN_TIP) != 0;
  const bool rate_scalers = (p->attributes & PLL_ATTRIB_RATE_SCALERS) != 0;

  for (unsigned int i = 0; i < count; ++i)
  {
    const pll_operation_t & o = ops_in[i];
    const bool bad =
        o.parent_clv_index >= p->nodes || o.parent_clv_index < c->clv_first ||
        o.child1_clv_index >= p->nodes || o.child2_clv_index >= p->nodes ||
        o.child1_matrix_index >= p->prob_matrices || o.child2_matrix_index >= p->prob_matrices ||
        o.parent_scaler_index >= (int)p->scale_buffers || o.child1_scaler_index >= (int)p->scale_buffers ||
        o.child2_scaler_index >= (int)p->scale_buffers;
    if (bad)
    {
      pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Operation %u references an invalid buffer.", i);
      return false;
    }
  }
  if (!dev_sync_inputs(c)) return false;

  const unsigned int K = p->rate_cats;
  if (p->states == 4 && (K == 1 || K == 2 || K == 4 || K == 8))
  {
    const pll_operation_t * ops = schedule_ops(c, ops_in, count);
    const pll_operation_t * d_ops =
        (const pll_operation_t *)dev_stage(c, ops, sizeof(pll_operation_t) * count);
    if (!d_ops) return false;
    TravArgs a;
    a.ops = d_ops;
    a.nops = count;
    a.clv = c->d_clv;
    a.clv_stride = c->clv_stride;
    a.clv_first = c->clv_first;
    a.scaler = c->d_scaler;
    a.scaler_stride = c->scaler_stride;
    a.pmat_il = c->d_pmat_il;
    a.tipchars = c->d_tipchars;
    a.tip_stride = c->tip_stride;
    a.tips_pattern = pattern_tip ? p->tips : 0;
    a.sites = c->sites_alloc;
    switch (K)
    {
      case 1: return launch_trav_dna<1>(c, a, rate_scalers);
      case 2: return launch_trav_dna<2>(c, a, rate_scalers);
      case 4: return launch_trav_dna<4>(c, a, rate_scalers);
      default: return launch_trav_dna<8>(c, a, rate_scalers);
    }
  }

  /* generic: one launch per op, caller's order */
  if (K > GEN_THREADS)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "rate_cats > %d is not supported.", GEN_THREADS);
    return false;
  }
  for (unsigned int i = 0; i < count; ++i)
  {
    const pll_operation_t & o = ops_in[i];
    GenOpArgs a;
    memset(&a, 0, sizeof(a));
    a.parent = clv_dev(c, o.parent_clv_index);
    if (pattern_tip && o.child1_clv_index < p->tips)
      a.ltip = c->d_tipchars + (size_t)o.child1_clv_index * c->tip_stride;
    else
      a.left = clv_dev(c, o.child1_clv_index);
    if (pattern_tip && o.child2_clv_index < p->tips)
      a.rtip = c->d_tipchars + (size_t)o.child2_clv_index * c->tip_stride;
    else
      a.right = clv_dev(c, o.child2_clv_index);
    a.lmat = c->d_pmat + (size_t)o.child1_matrix_index * c->pmat_elems;
    a.rmat = c->d_pmat + (size_t)o.child2_matrix_index * c->pmat_elems;
    a.pscaler = scaler_dev(c, o.parent_scaler_index);
    a.lscaler = scaler_dev(c, o.child1_scaler_index);
    a.rscaler = scaler_dev(c, o.child2_scaler_index);
    a.tipmap = c->d_tipmap;
    a.sites = c->sites_alloc;
    a.states = p->states;
    a.sp = p->states_padded;
    a.K = K;
    a.rate_scalers = rate_scalers;
    a.is4 = p->states == 4;
    const unsigned int spb = GEN_THREADS / K;
    const unsigned int blocks = (c->sites_alloc + spb - 1) / spb;
    partial_generic_kernel<<<blocks, GEN_THREADS, 0, c->stream>>>(a);
    ++g_kernel_launches;
    if (!pllb200_cuda_ok(cudaGetLastError(), "partial_generic_kernel launch")) return false;
  }
  return true;
}
