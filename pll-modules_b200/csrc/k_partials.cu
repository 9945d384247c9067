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
 * The 4-state site-column kernel itself lives in k_traversal_dna.cu.
 */
#include "internal.h"
#include "devutil.cuh"
#include "trav_schedule.h"

#include <algorithm>

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

int g_opt_generic_dmma = 1; /* 17..64 states (not 20): per-operation DMMA kernel instead of the scalar one */

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
/* 17..64 states except 20 (codon models: 61): one operation per launch on the FP64 tensor cores */
/* ------------------------------------------------------------------------------------------ */
/* X[site][i] = sum_j C[site][j] P[i][j] per rate category is a dense [sites x s] x [s x s] product
 * (core_partials.c:612-764 computes it with s^2 scalar FMAs per site and category): DMMA tiles
 * mma.sync.m8n8k4.f64 with M = 8 sites, N = 8 parent states (ceil(s/8) tiles), K = 4 child states (sp/4 steps).
 * A fragment (lane = 4 row + col): 256-bit pieces child[site0 + row][16 g + 4 col .. + 3], all requested before the
 * first mma; k-step 4 g + j takes component j (the K dimension is permuted, the staged matrix columns follow:
 * pd_col); B fragment from shared memory (row pitch 68 doubles: the 16 lanes of a half warp hit 16 different bank
 * pairs); D fragment: parent states 8 nt + 2 col, +1 of site row.  A CTA stages both matrices of ONE category (70 KB) and runs its 64 tiles of 8 sites through
 * them, category after category; a warp always owns the same tiles, so the per-site maximum over the categories
 * accumulates in shared memory without atomics, and the rare rescaling (all K s entries of a site < 2^-256:
 * core_partials.c:729-763) is a second pass of the warp over what it stored.  Tip children (tip patterns) are
 * sums of matrix columns over the set bits of the code's state mask.  Per-rate scalers: generic kernel. */
#define PD_WARPS 8
#define PD_PITCH 68
#define PD_TILES 64 /* tiles of 8 sites per CTA */
#define PD_SMEM (2 * 64 * PD_PITCH * sizeof(double) + PD_TILES * 8 * sizeof(int) + 2 * 64 * sizeof(double))

__device__ __forceinline__ void pd_dmma(double & c0, double & c1, double a, double b)
{
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

/* position of matrix column j in the staged image */
__device__ __forceinline__ unsigned int pd_col(unsigned int j) { return (j & ~15u) | ((j & 3u) << 2) | ((j >> 2) & 3u); }

__global__ void __launch_bounds__(PD_WARPS * 32) partial_dmma_kernel(GenOpArgs a, unsigned int tiles_per_cta)
{
  extern __shared__ __align__(16) double pd_sm[];
  double * sM[2] = {pd_sm, pd_sm + 64 * PD_PITCH};
  int * s_max = reinterpret_cast<int *>(pd_sm + 2 * 64 * PD_PITCH);
  double * s_rs = reinterpret_cast<double *>(s_max + PD_TILES * 8); /* [2][64] row sums: a fully ambiguous tip */
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row = lane >> 2, col = lane & 3;
  const unsigned int s = a.states, sp = a.sp, K = a.K;
  const unsigned int NT = (s + 7) / 8;
  const unsigned long long tile0 = (unsigned long long)blockIdx.x * tiles_per_cta;
  const bool tip[2] = {a.ltip != nullptr, a.rtip != nullptr};
  const bool tiptip = tip[0] && tip[1];
  const unsigned long long full_mask = s >= 64 ? ~0ull : (1ull << s) - 1ull;
  for (int i = threadIdx.x; i < PD_TILES * 8; i += blockDim.x) s_max[i] = -2147483647 - 1;

  for (unsigned int k = 0; k < K; ++k)
  {
    __syncthreads(); /* everybody is done with the previous category's matrices */
    for (unsigned int idx = threadIdx.x; idx < 2 * 64 * 64; idx += blockDim.x)
    {
      const unsigned int q = idx >> 12, i = (idx >> 6) & 63u, j = idx & 63u;
      const double * m = q ? a.rmat : a.lmat;
      /* columns are stored with the 4 x 4 blocks of every 16 transposed (see the A fragment below) */
      sM[q][i * PD_PITCH + pd_col(j)] = (i < s && j < s) ? __ldg(m + ((size_t)k * s + i) * sp + j) : 0.0;
    }
    __syncthreads();
    if ((tip[0] || tip[1]) && threadIdx.x < 128)
    {
      /* gaps / unknown characters carry every state: their masked sum is the row sum, added in state order like
       * the loop below — one lookup instead of s iterations for the whole warp */
      const double * Mr = sM[threadIdx.x >> 6] + (threadIdx.x & 63) * PD_PITCH;
      double v = 0.0;
      for (unsigned int j = 0; j < s; ++j) v += Mr[pd_col(j)];
      s_rs[threadIdx.x] = v;
    }
    if (tip[0] || tip[1]) __syncthreads();
    for (unsigned int t = warp; t < tiles_per_cta; t += PD_WARPS)
    {
      const unsigned long long site = (tile0 + t) * 8 + row;
      if ((tile0 + t) * 8 >= a.sites) break; /* warp-uniform */
      const bool act = site < a.sites;
      double D[2][8][2];
#pragma unroll
      for (int q = 0; q < 2; ++q)
      {
        const double * M = sM[q];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) D[q][nt][0] = D[q][nt][1] = 0.0;
        if (tip[q])
        {
          unsigned long long mask = 0;
          if (act) mask = a.tipmap[(q ? a.rtip : a.ltip)[site]];
          if (mask == full_mask)
          {
#pragma unroll
            for (int nt = 0; nt < 8; ++nt)
              if ((unsigned int)nt < NT)
              {
                D[q][nt][0] = s_rs[q * 64 + nt * 8 + 2 * col];
                D[q][nt][1] = s_rs[q * 64 + nt * 8 + 2 * col + 1];
              }
            mask = 0;
          }
          while (mask)
          {
            const unsigned int j = (unsigned int)__ffsll((long long)mask) - 1u;
            mask &= mask - 1;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt)
              if ((unsigned int)nt < NT)
              {
                D[q][nt][0] += M[(nt * 8 + 2 * col) * PD_PITCH + pd_col(j)];
                D[q][nt][1] += M[(nt * 8 + 2 * col + 1) * PD_PITCH + pd_col(j)];
              }
          }
        }
        else
        {
          /* the K dimension of an MMA may be permuted as long as B follows: lane (row, col) loads 256-bit pieces
           * child[16 g + 4 col .. + 3] (a quad of lanes = one 128-byte line of its site), k-step 4 g + j takes
           * component j: child state e = 16 g + 4 col + j, whose matrix column sits at 16 g + 4 j + col */
          const double * cv = (q ? a.right : a.left) + ((size_t)site * K + k) * sp + 4 * col;
          d4 piece[4];
#pragma unroll
          for (int g = 0; g < 4; ++g)
          {
            piece[g] = d4{0.0, 0.0, 0.0, 0.0};
            if (act && 16u * g + 4u * col < sp) piece[g] = ldg256(cv + 16 * g);
          }
#pragma unroll
          for (int g = 0; g < 4; ++g)
          {
            if (16u * g >= sp) break; /* warp-uniform */
            const unsigned int e0 = 16u * g + 4u * col;
            /* padding is never used (it may hold anything) */
            const double av[4] = {e0 < s ? piece[g].x : 0.0, e0 + 1 < s ? piece[g].y : 0.0, e0 + 2 < s ? piece[g].z : 0.0,
                                  e0 + 3 < s ? piece[g].w : 0.0};
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
              for (int nt = 0; nt < 8; ++nt)
                if ((unsigned int)nt < NT)
                  pd_dmma(D[q][nt][0], D[q][nt][1], av[j], M[(nt * 8 + row) * PD_PITCH + 16 * g + 4 * j + col]);
          }
        }
      }
      int m = -2147483647 - 1;
      double * dst = a.parent + ((size_t)site * K + k) * sp;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
        if ((unsigned int)nt < NT)
        {
          const double o0 = D[0][nt][0] * D[1][nt][0], o1 = D[0][nt][1] * D[1][nt][1];
          const unsigned int i0 = nt * 8 + 2 * col;
          if (i0 < s) m = max(m, __double2hiint(o0));
          if (i0 + 1 < s) m = max(m, __double2hiint(o1));
          if (act && i0 < sp) *reinterpret_cast<double2 *>(dst + i0) = make_double2(o0, o1); /* sp is a multiple of 4 */
        }
      m = max(m, __shfl_xor_sync(0xffffffffu, m, 1));
      m = max(m, __shfl_xor_sync(0xffffffffu, m, 2));
      if (col == 0) s_max[t * 8 + row] = max(s_max[t * 8 + row], m);
    }
  }
  if (!a.pscaler) return;
  __syncwarp();
  /* scaler counts and the rare rescaling pass: the warp's own tiles */
  for (unsigned int t = warp; t < tiles_per_cta; t += PD_WARPS)
  {
    const unsigned long long site = (tile0 + t) * 8 + row;
    if ((tile0 + t) * 8 >= a.sites) break;
    const bool act = site < a.sites;
    const bool scale = act && !tiptip && s_max[t * 8 + row] < 0x2FF00000;
    if (act && col == 0)
    {
      unsigned int v = scale ? 1u : 0u;
      if (a.lscaler) v += a.lscaler[site];
      if (a.rscaler) v += a.rscaler[site];
      a.pscaler[site] = v;
    }
    if (__any_sync(0xffffffffu, scale))
    {
      __syncwarp(); /* the stores of the other lanes of this warp */
      if (scale)
      {
        double * rowp = a.parent + (size_t)site * K * sp;
        for (unsigned int e = col; e < K * sp; e += 4) rowp[e] *= SCALE_FACTOR_D;
      }
    }
  }
}

/* ---- host side of the 17..64-state tensor-core kernel ------------------------------------------------ */
bool generic_dmma_usable(const PartCtx * c)
{
  const pll_partition_t * p = &c->pub;
  return g_opt_generic_dmma && p->states > 16 && p->states <= 64 && p->states_padded <= 64 &&
         !(p->attributes & PLL_ATTRIB_RATE_SCALERS);
}

static bool launch_partial_dmma(PartCtx * c, const GenOpArgs & a, cudaStream_t stream)
{
  static bool attr_set = false;
  if (!attr_set)
  {
    if (!pllb200_cuda_ok(cudaFuncSetAttribute(partial_dmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PD_SMEM),
                         "partial_dmma_kernel smem"))
      return false;
    attr_set = true;
  }
  /* tiles of 8 sites per CTA: a multiple of the warp count, sized so that the grid is about one resident wave
   * (2 CTAs per SM); a CTA re-stages the matrices of every category, so more than that only costs L2 reads */
  const unsigned int tiles = (a.sites + 7) / 8;
  unsigned int tpc = (tiles + 2 * c->sm_count - 1) / (2 * c->sm_count);
  tpc = (tpc + PD_WARPS - 1) / PD_WARPS * PD_WARPS;
  if (tpc > PD_TILES) tpc = PD_TILES;
  if (tpc < PD_WARPS) tpc = PD_WARPS;
  partial_dmma_kernel<<<(tiles + tpc - 1) / tpc, PD_WARPS * 32, PD_SMEM, stream>>>(a, tpc);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "partial_dmma_kernel launch");
}

/* The sumtable of an edge (A.6 of SURVEY.md, core_derivatives.c:116-640) IS a partial-likelihood operation whose
 * "P-matrices" are the eigenvector images  L[j][i] = pi_i Vinv[i][j]  and  R[j][i] = V[j][i], whose parent buffer is
 * the sumtable, and which never scales (the trick of launch_sumtable_aa, k_traversal_aa.cu).  The scalar kernel
 * took 2.2 ms per edge at 61 states x 20 k sites = 81 % of a smoothing pass
 * (profiles/r02h_launches_blo_codon_100x20k.txt). */
struct SumGenIdx
{
  unsigned int v[32];
};

__global__ void sumgen_kernel(double * __restrict__ out, const double * __restrict__ eigen, unsigned long long eigen_stride,
                              unsigned int K, unsigned int s, unsigned int sp, SumGenIdx idx)
{
  const unsigned int k = blockIdx.x;
  const double * blk = eigen + (unsigned long long)idx.v[k] * eigen_stride;
  const double * ev = blk, * iev = blk + s * sp, * fr = blk + 2 * s * sp + sp;
  double * L = out + (size_t)k * s * sp, * R = out + ((size_t)K + k) * s * sp;
  for (unsigned int t = threadIdx.x; t < s * sp; t += blockDim.x)
  {
    const unsigned int j = t / sp, i = t % sp;
    L[t] = i < s ? fr[i] * iev[i * sp + j] : 0.0;
    R[t] = i < s ? ev[j * sp + i] : 0.0;
  }
}

bool launch_sumtable_dmma(PartCtx * c, bool first_is_tip, unsigned int first, unsigned int second,
                          const unsigned int * params_indices, double * d_sumtable)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats, s = p->states, sp = p->states_padded;
  if (K > 32) return false;
  const size_t doubles = (size_t)2 * K * s * sp;
  if (!c->d_sumgen &&
      !pllb200_cuda_ok(cudaMalloc(&c->d_sumgen, doubles * sizeof(double)), "sumtable eigenvector images"))
    return false;
  std::vector<unsigned long long> key(K);
  for (unsigned int k = 0; k < K; ++k)
    key[k] = ((unsigned long long)params_indices[k] << 32) | c->eigen_version[params_indices[k]];
  if (key != c->sumgen_key)
  {
    SumGenIdx idx;
    memset(&idx, 0, sizeof(idx));
    for (unsigned int k = 0; k < K; ++k) idx.v[k] = params_indices[k];
    sumgen_kernel<<<K, 256, 0, xstream(c)>>>(c->d_sumgen, c->d_eigen, c->eigen_stride, K, s, sp, idx);
    ++g_kernel_launches;
    if (!pllb200_cuda_ok(cudaGetLastError(), "sumgen_kernel launch")) return false;
    c->sumgen_key = key;
  }
  GenOpArgs a;
  memset(&a, 0, sizeof(a));
  a.parent = d_sumtable;
  if (first_is_tip)
    a.ltip = c->d_tipchars + (size_t)first * c->tip_stride;
  else
    a.left = clv_dev(c, first);
  a.right = clv_dev(c, second);
  a.lmat = c->d_sumgen;
  a.rmat = c->d_sumgen + (size_t)K * s * sp;
  a.tipmap = c->d_tipmap;
  a.sites = c->sites_alloc;
  a.states = s;
  a.sp = sp;
  a.K = K;
  return launch_partial_dmma(c, a, xstream(c));
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
    /* Sethi-Ullman order (trav_schedule.cpp): what the forwarding + stash kernel wants; the other 4-state
     * variants take the same order (it is depth-first too) */
    const pll_operation_t * ops = ops_in;
    if (g_opt_trav_reorder && trav_order_ops(ops_in, count, p->nodes, true, c->op_scratch)) ops = c->op_scratch.data();
    return launch_trav_dna(c, ops, count);
  }

  if (p->states == 20 && !rate_scalers && K <= 8 && c->d_pmat_il)
    return launch_trav_aa(c, schedule_ops(c, ops_in, count), count); /* FP64 tensor-core path */

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
    if (generic_dmma_usable(c) && (!pattern_tip || c->d_tipmap))
    {
      if (!launch_partial_dmma(c, a, c->stream)) return false;
      continue;
    }
    const unsigned int spb = GEN_THREADS / K;
    const unsigned int blocks = (c->sites_alloc + spb - 1) / spb;
    partial_generic_kernel<<<blocks, GEN_THREADS, 0, c->stream>>>(a);
    ++g_kernel_launches;
    if (!pllb200_cuda_ok(cudaGetLastError(), "partial_generic_kernel launch")) return false;
  }
  return true;
}
