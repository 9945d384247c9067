/*
 * k_pmatrix.cu — batched transition-probability matrices.
 *
 * Replaces pll_core_update_pmatrix (libs/libpll/src/core_pmatrix.c:24-258 and its SSE/AVX/AVX2
 * twins).  One CTA per (edge, rate category): P = I + V^-1 * diag(expm1(lambda_j * r_k * t / (1-pinv))) * V
 * with the reference's conventions — the (1-pinv) divisor only when pinv > PLL_MISC_EPSILON
 * (core_pmatrix.c:210), t == 0 gives the identity (:238-245), expm1 + identity for accuracy at
 * small t (:201-206).  All `count` edges of a call go out in ONE launch (the reference's callers
 * hand over one matrix per call, treeinfo.c:854 — our treeinfo batches them).
 *
 * Roofline: irrelevant (count*K*s*sp*8 bytes, ~1 MB for 2000 DNA edges); latency is what matters,
 * hence one launch and kernel-parameter passing of the per-category model scalars.
 */
#include "internal.h"
#include "devutil.cuh"

#define PM_MAXK 32

struct PmatArgs
{
  unsigned int params_idx[PM_MAXK];
  double rates[PM_MAXK];
  double pinv[PM_MAXK];
  /* a single matrix travels in the kernel parameters: index when `matrix_indices` is null, length when
   * `branch_lengths` is null */
  unsigned int one_index;
  double one_length;
  /* 20 states with tip patterns: also emit, per tip code c < 24, sum_{j in mask(c)} P[i][j] (internal.h) */
  double * tiptab;
  const unsigned long long * tipmap;
  /* 4 states: also emit the column table of k_traversal_dna4.cu, [m][A C G T N][k][row i] */
  double * cols;
};

#define PM_TIP_CODES 24

/* rows of the tip table of matrix (m, k) from the finished matrix `out` (written by this CTA) */
__device__ __forceinline__ void pmatrix_tip_rows(const double * out, double * tiptab, const unsigned long long * tipmap,
                                                 unsigned int m, unsigned int k, unsigned int K, unsigned int sp)
{
  double * tab = tiptab + (size_t)m * K * (PM_TIP_CODES * 20);
  for (unsigned int idx = threadIdx.x; idx < PM_TIP_CODES * 20; idx += blockDim.x)
  {
    const unsigned int code = idx / 20, i = idx % 20;
    const unsigned long long mask = tipmap[code];
    double v = 0.0;
    for (unsigned int j = 0; j < 20; ++j)
      if ((mask >> j) & 1ull) v += out[i * sp + j];
    tab[(code * K + k) * 20 + i] = v;
  }
}

/* column table of matrix (m, k): what a DNA tip code contributes to parent row i — the column of a
 * single base, or for N the row sum added in state order like the masked sums of core_partials.c:290-351 */
__device__ __forceinline__ void pmatrix_dna_cols(const double * out, double * cols, unsigned int m, unsigned int k,
                                                 unsigned int K, unsigned int sp)
{
  if (threadIdx.x < 20)
  {
    const unsigned int c = threadIdx.x >> 2, i = threadIdx.x & 3u;
    double v;
    if (c < 4)
      v = out[i * sp + c];
    else
      v = ((out[i * sp] + out[i * sp + 1]) + out[i * sp + 2]) + out[i * sp + 3];
    /* odd columns keep rows 2-3 before rows 0-1 (shared-memory banks of the traversal kernel's lookups:
     * tip_term, k_traversal_dna4.cu) */
    cols[(((size_t)m * 5 + c) * K + k) * 4 + ((c & 1u) ? (i ^ 2u) : i)] = v;
  }
}

__device__ __forceinline__ void pmatrix_body(double * __restrict__ pmat, double * __restrict__ pmat_il,
                                             const double * __restrict__ eigen, size_t eigen_stride,
                                             unsigned int m, double t, unsigned int k, unsigned int states,
                                             unsigned int sp, unsigned int K, const PmatArgs & a)
{
  extern __shared__ double sm[]; /* expd[sp] | inv[s*sp] | ev[s*sp] */
  double * expd = sm;
  double * s_inv = sm + sp;
  double * s_ev = s_inv + states * sp;
  const double * blk = eigen + (size_t)a.params_idx[k] * eigen_stride;
  const double * g_ev = blk;
  const double * g_inv = blk + states * sp;
  const double * g_val = blk + 2 * states * sp;

  for (unsigned int i = threadIdx.x; i < states * sp; i += blockDim.x)
  {
    s_inv[i] = g_inv[i];
    s_ev[i] = g_ev[i];
  }
  if (threadIdx.x < states)
  {
    const double pinv = a.pinv[k];
    double x = g_val[threadIdx.x] * a.rates[k] * t;
    if (pinv > PLL_MISC_EPSILON) x = g_val[threadIdx.x] * a.rates[k] * t / (1.0 - pinv);
    expd[threadIdx.x] = expm1(x);
  }
  __syncthreads();

  double * out = pmat + ((size_t)m * K + k) * states * sp;
  for (unsigned int idx = threadIdx.x; idx < states * sp; idx += blockDim.x)
  {
    const unsigned int i = idx / sp, j = idx % sp;
    double v = 0.0;
    if (j < states)
    {
      v = (i == j) ? 1.0 : 0.0;
      if (t > 0.0)
        for (unsigned int q = 0; q < states; ++q) v += (s_inv[i * sp + q] * expd[q]) * s_ev[q * sp + j];
    }
    out[idx] = v;
    if (pmat_il && states == 4) /* 4-state: [m][i][k][j] */
      pmat_il[(size_t)m * K * 16 + (i * K + k) * 4 + j] = v;
    if (pmat_il && states == 20)
    {
      /* m8n8k4 B fragment (col-major 4x8): lane l holds B[row l%4][col l/4] = P[i][j] with i the state of
       * column l/4 of n-tile nt and j the state of k-slot (ks, l%4) (devutil.cuh); columns 4..7 of the third n-tile stay zero from the allocation-time memset */
      const unsigned int nt = aa_ntile_of_state(i), ks = aa_kstep_of_state(j), lane = (aa_ncol_of_state(i) << 2) | aa_col_of_state(j);
      pmat_il[(((size_t)m * K + k) * 3 + nt) * 160 + ks * 32 + lane] = v;
    }
  }
  if (a.tiptab && states == 20)
  {
    __syncthreads(); /* `out` complete and visible to the whole CTA */
    pmatrix_tip_rows(out, a.tiptab, a.tipmap, m, k, K, sp);
  }
  if (a.cols && states == 4)
  {
    __syncthreads();
    pmatrix_dna_cols(out, a.cols, m, k, K, sp);
  }
}

__global__ void pmatrix_kernel(double * __restrict__ pmat, double * __restrict__ pmat_il,
                               const double * __restrict__ eigen, size_t eigen_stride,
                               const unsigned int * __restrict__ matrix_indices,
                               const double * __restrict__ branch_lengths, unsigned int states,
                               unsigned int sp, unsigned int K, PmatArgs a)
{
  const unsigned int e = blockIdx.x / K;
  const unsigned int k = blockIdx.x % K;
  const unsigned int m = matrix_indices ? matrix_indices[e] : a.one_index;
  const double t = branch_lengths ? branch_lengths[e] : a.one_length;
  pmatrix_body(pmat, pmat_il, eigen, eigen_stride, m, t, k, states, sp, K, a);
}

/* one matrix of each of many partitions in one launch (blockIdx.y = partition, blockIdx.x = category;
 * batch mode, internal.h) */
struct PmatBatchPart
{
  const unsigned int * idx; /* pool tokens -> device pointers at flush (first two members, device_ctx.cu); */
  const double * len;       /* null: one matrix, index and length in `a` */
  double * pmat;
  double * pmat_il;
  const double * eigen;
  unsigned long long eigen_stride;
  unsigned int states, sp, K, count;
  PmatArgs a;
};

__global__ void pmatrix_batch_kernel(const PmatBatchPart * __restrict__ parts)
{
  __shared__ PmatBatchPart sp_;
  for (unsigned int i = threadIdx.x; i < sizeof(PmatBatchPart) / 4; i += blockDim.x)
    reinterpret_cast<unsigned int *>(&sp_)[i] = reinterpret_cast<const unsigned int *>(parts + blockIdx.y)[i];
  __syncthreads();
  const unsigned int e = blockIdx.x / sp_.K, k = blockIdx.x % sp_.K;
  if (e >= sp_.count) return;
  const unsigned int m = sp_.idx ? sp_.idx[e] : sp_.a.one_index;
  const double t = sp_.len ? sp_.len[e] : sp_.a.one_length;
  pmatrix_body(sp_.pmat, sp_.pmat_il, sp_.eigen, sp_.eigen_stride, m, t, k, sp_.states, sp_.sp, sp_.K, sp_.a);
}

/* tip tables are used when the partition has them and every tip code fits */
static bool tiptab_usable(const PartCtx * c) { return c->d_tiptab && c->d_tipmap && c->pub.maxstates <= PM_TIP_CODES; }

/* one matrix whose branch length is already on the device (written by the queued Newton kernel) */
/* more than 48 KB of dynamic shared memory (more than 54 states) needs an opt-in per kernel */
static bool pmat_smem_optin(size_t smem)
{
  static size_t granted = 48 * 1024;
  if (smem <= granted) return true;
  if (!pllb200_cuda_ok(cudaFuncSetAttribute(pmatrix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                       "pmatrix_kernel shared memory") ||
      !pllb200_cuda_ok(cudaFuncSetAttribute(pmatrix_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                       "pmatrix_batch_kernel shared memory"))
    return false;
  granted = smem;
  return true;
}

bool launch_pmatrix_devlen(PartCtx * c, const unsigned int * params_indices, unsigned int matrix_index,
                           const double * d_length)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats, s = p->states, sp = p->states_padded;
  if (K > PM_MAXK || matrix_index >= p->prob_matrices)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid matrix index %u or rate_cats > %d.", matrix_index, PM_MAXK);
    return false;
  }
  if (!dev_sync_eigen(c, params_indices)) return false;
  PmatArgs a;
  for (unsigned int k = 0; k < K; ++k)
  {
    a.params_idx[k] = params_indices[k];
    a.rates[k] = p->rates[k];
    a.pinv[k] = p->prop_invar[params_indices[k]];
  }
  a.tiptab = nullptr;
  a.tipmap = nullptr;
  a.cols = c->d_pmat_cols;
  if (tiptab_usable(c))
  {
    if (!dev_sync_inputs(c)) return false;
    a.tiptab = c->d_tiptab;
    a.tipmap = c->d_tipmap;
  }
  a.one_index = matrix_index;
  a.one_length = 0.0;
  const unsigned int * d_idx = nullptr;
  unsigned int threads = ((s * sp + 31) / 32) * 32;
  if (threads > 512) threads = 512;
  const size_t smem = sizeof(double) * (sp + 2 * s * sp);
  if (!pmat_smem_optin(smem)) return false;
  pmatrix_kernel<<<K, threads, smem, c->stream>>>(c->d_pmat, c->d_pmat_il, c->d_eigen, c->eigen_stride, d_idx,
                                                  d_length, s, sp, K, a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "pmatrix_kernel launch");
}

/* rebuild the interleaved copies from the plain array (after P-matrices were uploaded from a
 * checkpoint, binary_io.cpp): same index maps as pmatrix_kernel */
__global__ void pmatrix_reinterleave_kernel(const double * __restrict__ pmat, double * __restrict__ pmat_il,
                                            unsigned int states, unsigned int sp, unsigned int K,
                                            double * tiptab, const unsigned long long * tipmap, double * cols)
{
  const unsigned int m = blockIdx.x / K, k = blockIdx.x % K;
  const double * in = pmat + ((size_t)m * K + k) * states * sp;
  for (unsigned int idx = threadIdx.x; idx < states * sp; idx += blockDim.x)
  {
    const unsigned int i = idx / sp, j = idx % sp;
    const double v = in[idx];
    if (states == 4)
      pmat_il[(size_t)m * K * 16 + (i * K + k) * 4 + j] = v;
    else if (states == 20)
    {
      const unsigned int nt = aa_ntile_of_state(i), ks = aa_kstep_of_state(j), lane = (aa_ncol_of_state(i) << 2) | aa_col_of_state(j);
      pmat_il[(((size_t)m * K + k) * 3 + nt) * 160 + ks * 32 + lane] = v;
    }
  }
  if (tiptab && states == 20) pmatrix_tip_rows(in, tiptab, tipmap, m, k, K, sp);
  if (cols && states == 4) pmatrix_dna_cols(in, cols, m, k, K, sp);
}

bool launch_pmatrix_reinterleave(PartCtx * c)
{
  pll_partition_t * p = &c->pub;
  if (!c->d_pmat_il || !p->prob_matrices || (p->states != 4 && p->states != 20)) return true;
  if (tiptab_usable(c) && !dev_sync_inputs(c)) return false; /* the tip map must be on the device */
  c->tiptab_stale = false;
  pmatrix_reinterleave_kernel<<<p->prob_matrices * p->rate_cats, 128, 0, xstream(c)>>>(
      c->d_pmat, c->d_pmat_il, p->states, p->states_padded, p->rate_cats, tiptab_usable(c) ? c->d_tiptab : nullptr,
      c->d_tipmap, c->d_pmat_cols);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "pmatrix_reinterleave_kernel launch");
}

bool launch_pmatrix(PartCtx * c, const unsigned int * params_indices,
                    const unsigned int * matrix_indices, const double * branch_lengths,
                    unsigned int count)
{
  pll_partition_t * p = &c->pub;
  const unsigned int K = p->rate_cats, s = p->states, sp = p->states_padded;
  if (K > PM_MAXK)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "rate_cats > %d is not supported.", PM_MAXK);
    return false;
  }
  PmatArgs a;
  for (unsigned int k = 0; k < K; ++k)
  {
    a.params_idx[k] = params_indices[k];
    a.rates[k] = p->rates[k];
    a.pinv[k] = p->prop_invar[params_indices[k]];
  }
  a.tiptab = nullptr;
  a.tipmap = nullptr;
  a.cols = c->d_pmat_cols;
  if (tiptab_usable(c))
  {
    if (!dev_sync_inputs(c)) return false;
    a.tiptab = c->d_tiptab;
    a.tipmap = c->d_tipmap;
  }
  for (unsigned int i = 0; i < count; ++i)
  {
    if (matrix_indices[i] >= p->prob_matrices + c->scratch_mats || !(branch_lengths[i] >= 0.0))
    {
      pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid matrix index %u or branch length %g.",
                        matrix_indices[i], branch_lengths[i]);
      return false;
    }
  }
  const unsigned int * d_idx = nullptr;
  const double * d_len = nullptr;
  a.one_index = count ? matrix_indices[0] : 0;
  a.one_length = count ? branch_lengths[0] : 0.0;
  unsigned int threads = ((s * sp + 31) / 32) * 32;
  if (threads > 512) threads = 512;
  const size_t smem = sizeof(double) * (sp + 2 * s * sp);
  if (!pmat_smem_optin(smem)) return false;
  if (batch_collects(BK_PMAT))
  {
    PmatBatchPart b;
    memset(&b, 0, sizeof(b));
    b.count = count;
    if (count != 1)
    {
      b.idx = (const unsigned int *)batch_pool_add(matrix_indices, sizeof(unsigned int) * count);
      b.len = (const double *)batch_pool_add(branch_lengths, sizeof(double) * count);
    }
    b.pmat = c->d_pmat;
    b.pmat_il = c->d_pmat_il;
    b.eigen = c->d_eigen;
    b.eigen_stride = c->eigen_stride;
    b.states = s;
    b.sp = sp;
    b.K = K;
    b.a = a;
    batch_collect(BK_PMAT, &b, sizeof(b), count * K, smem);
    return true;
  }
  if (count != 1)
  {
    d_idx = (const unsigned int *)dev_stage_small(c, matrix_indices, sizeof(unsigned int) * count);
    d_len = (const double *)dev_stage_small(c, branch_lengths, sizeof(double) * count);
    if (!d_idx || !d_len) return false;
  }
  pmatrix_kernel<<<count * K, threads, smem, xstream(c)>>>(c->d_pmat, c->d_pmat_il, c->d_eigen,
                                                           c->eigen_stride, d_idx, d_len, s, sp, K, a);
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "pmatrix_kernel launch");
}

bool flush_pmat(PartCtx * lead, const void * d_args, const BatchKind & k)
{
  /* max_blocks = largest count * K; the thread count is the largest any member needs */
  unsigned int threads = 32;
  const PmatBatchPart * h = reinterpret_cast<const PmatBatchPart *>(k.bytes.data());
  for (unsigned int i = 0; i < k.count; ++i)
  {
    unsigned int t = ((h[i].states * h[i].sp + 31) / 32) * 32;
    if (t > 512) t = 512;
    threads = t > threads ? t : threads;
  }
  pmatrix_batch_kernel<<<dim3(k.max_blocks, k.count), threads, k.smem, lead->stream>>>(
      static_cast<const PmatBatchPart *>(d_args));
  ++g_kernel_launches;
  return pllb200_cuda_ok(cudaGetLastError(), "batched pmatrix kernel launch");
}
