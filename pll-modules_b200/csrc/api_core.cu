/*
 * api_core.cu — the six libpll entry points of the hot path, as thin dispatchers onto the kernels.
 *
 *   pll_update_prob_matrices            models.c:412      -> k_pmatrix.cu
 *   pll_update_partials                 partials.c:237    -> k_partials.cu
 *   pll_compute_edge_loglikelihood      likelihood.c:586  -> k_likelihood.cu
 *   pll_compute_root_loglikelihood      likelihood.c:122  -> k_likelihood.cu
 *   pll_update_sumtable                 derivatives.c:239 -> k_derivatives.cu
 *   pll_compute_likelihood_derivatives  derivatives.c:333 -> k_derivatives.cu
 *
 * Calls that return no scalar are queued on the partition's stream and return immediately; calls
 * that return a scalar wait for the stream and read the value from mapped pinned memory.
 * Error convention as the reference: PLL_FAILURE / -INFINITY plus pll_errno + pll_errmsg.
 */
#include "internal.h"

#include <vector>

#include <cmath>
#include <mutex>

static bool enter(pll_partition_t * p)
{
  if (!p || ctx_of(p)->magic != PLLB200_MAGIC)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Not a libpll_b200 partition.");
    return false;
  }
  return pllb200_cuda_ok(cudaSetDevice(ctx_of(p)->device), "cudaSetDevice");
}

extern "C" int pll_update_prob_matrices(pll_partition_t * partition,
                                        const unsigned int * params_indices,
                                        const unsigned int * matrix_indices,
                                        const double * branch_lengths, unsigned int count)
{
  if (!enter(partition)) return PLL_FAILURE;
  PartCtx * c = ctx_of(partition);
  if (!count) return PLL_SUCCESS;
  if (!dev_sync_eigen(c, params_indices)) return PLL_FAILURE;
  return launch_pmatrix(c, params_indices, matrix_indices, branch_lengths, count) ? PLL_SUCCESS
                                                                                  : PLL_FAILURE;
}

extern "C" void pll_update_partials(pll_partition_t * partition,
                                    const pll_operation_t * operations, unsigned int count)
{
  if (!enter(partition)) return;
  launch_partials(ctx_of(partition), operations, count);
}

static double * persite_buffer(PartCtx * c)
{
  if (!c->d_persite)
    if (!pllb200_cuda_ok(cudaMalloc(&c->d_persite, sizeof(double) * c->pub.sites), "persite alloc"))
      return nullptr;
  return c->d_persite;
}

static double finish_scalar(PartCtx * c, int slot, double * persite_lnl, double * d_persite)
{
  if (persite_lnl)
  {
    if (!pllb200_cuda_ok(cudaMemcpyAsync(persite_lnl, d_persite, sizeof(double) * c->pub.sites,
                                         cudaMemcpyDeviceToHost, c->stream), "persite copy"))
      return -INFINITY;
    g_d2h_bytes += sizeof(double) * c->pub.sites;
  }
  /* with a per-site copy queued behind the kernel the stream has to drain; otherwise poll the tag */
  if (persite_lnl ? !dev_flush(c) : !(dev_wait_results(c) && dev_wait_slot(c, slot))) return -INFINITY;
  g_d2h_bytes += sizeof(double);
  return c->h_result[slot * 4];
}

extern "C" double pll_compute_edge_loglikelihood(pll_partition_t * partition,
                                                 unsigned int parent_clv_index,
                                                 int parent_scaler_index,
                                                 unsigned int child_clv_index,
                                                 int child_scaler_index,
                                                 unsigned int matrix_index,
                                                 const unsigned int * freqs_indices,
                                                 double * persite_lnl)
{
  if (!enter(partition)) return -INFINITY;
  PartCtx * c = ctx_of(partition);
  double * d_ps = nullptr;
  if (persite_lnl && !(d_ps = persite_buffer(c))) return -INFINITY;
  const int slot = dev_alloc_slot(c);
  if (!launch_edge_logl(c, parent_clv_index, parent_scaler_index, child_clv_index,
                        child_scaler_index, matrix_index, freqs_indices, d_ps, slot))
    return -INFINITY;
  return finish_scalar(c, slot, persite_lnl, d_ps);
}

/* likelihood.c:639-823: marginal ancestral state probabilities of an inner node, [site][state] */
extern "C" int pll_compute_node_ancestral(pll_partition_t * partition, unsigned int node_clv_index,
                                          int node_scaler_index, unsigned int other_clv_index,
                                          int other_scaler_index, unsigned int matrix_index,
                                          const unsigned int * freqs_indices, double * ancestral)
{
  if (!enter(partition)) return PLL_FAILURE;
  if (!ancestral)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Parameter value is NULL!");
    return PLL_FAILURE;
  }
  PartCtx * c = ctx_of(partition);
  const size_t bytes = sizeof(double) * partition->sites * partition->states;
  if (!c->d_ancestral && !pllb200_cuda_ok(cudaMalloc(&c->d_ancestral, bytes), "ancestral buffer"))
    return PLL_FAILURE;
  if (!launch_ancestral(c, node_clv_index, node_scaler_index, other_clv_index, other_scaler_index,
                        matrix_index, freqs_indices, c->d_ancestral))
    return PLL_FAILURE;
  if (!pllb200_cuda_ok(cudaMemcpyAsync(ancestral, c->d_ancestral, bytes, cudaMemcpyDeviceToHost, c->stream),
                       "ancestral D2H"))
    return PLL_FAILURE;
  g_d2h_bytes += bytes;
  return dev_flush(c) ? PLL_SUCCESS : PLL_FAILURE;
}

/* the caller's scratch buffers (temporary CLV, scaler, identity matrix) are not needed on the device */
extern "C" int pll_compute_node_ancestral_extbuf(pll_partition_t * partition, unsigned int node_clv_index,
                                                 int node_scaler_index, unsigned int other_clv_index,
                                                 int other_scaler_index, unsigned int pmatrix_index,
                                                 const unsigned int * freqs_indices, double * ancestral,
                                                 double * temp_clv, unsigned int * temp_scaler,
                                                 double * ident_pmat)
{
  (void)temp_clv;
  (void)temp_scaler;
  (void)ident_pmat;
  return pll_compute_node_ancestral(partition, node_clv_index, node_scaler_index, other_clv_index,
                                    other_scaler_index, pmatrix_index, freqs_indices, ancestral);
}

extern "C" int pllb200_edge_loglikelihood_async(pll_partition_t * partition,
                                                unsigned int parent_clv_index,
                                                int parent_scaler_index,
                                                unsigned int child_clv_index,
                                                int child_scaler_index, unsigned int matrix_index,
                                                const unsigned int * freqs_indices, double * result)
{
  if (!enter(partition)) return PLL_FAILURE;
  PartCtx * c = ctx_of(partition);
  const int slot = dev_alloc_slot(c);
  if (!launch_edge_logl(c, parent_clv_index, parent_scaler_index, child_clv_index,
                        child_scaler_index, matrix_index, freqs_indices, nullptr, slot))
    return PLL_FAILURE;
  c->pending.push_back(PendingResult{result, slot, 0});
  return PLL_SUCCESS;
}

extern "C" double pll_compute_root_loglikelihood(pll_partition_t * partition,
                                                 unsigned int clv_index, int scaler_index,
                                                 const unsigned int * freqs_indices,
                                                 double * persite_lnl)
{
  if (!enter(partition)) return -INFINITY;
  PartCtx * c = ctx_of(partition);
  double * d_ps = nullptr;
  if (persite_lnl && !(d_ps = persite_buffer(c))) return -INFINITY;
  const int slot = dev_alloc_slot(c);
  if (!launch_root_logl(c, clv_index, scaler_index, freqs_indices, d_ps, slot)) return -INFINITY;
  return finish_scalar(c, slot, persite_lnl, d_ps);
}

/* device shadow of a caller-owned sumtable buffer.  Keys are also kept in a
 * process-wide registry so that pll_aligned_free(key) gives the HBM back. */
static std::mutex g_key_mutex;
static std::unordered_map<const void *, PartCtx *> g_key_owner;

void sumtable_forget_key(const void * key)
{
  std::lock_guard<std::mutex> lock(g_key_mutex);
  if (g_key_owner.empty()) return;
  auto it = g_key_owner.find(key);
  if (it == g_key_owner.end()) return;
  PartCtx * c = it->second;
  g_key_owner.erase(it);
  auto sh = c->sumtables.find(key);
  if (sh != c->sumtables.end())
  {
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaFree(sh->second);
    c->sumtables.erase(sh);
    c->sumtable_stamp.erase(key);
  }
}

/* additive: release the device shadow of a sumtable buffer that is about to be freed with plain free() */
extern "C" void pllb200_forget_sumtable(const void * sumtable_key) { sumtable_forget_key(sumtable_key); }

void sumtable_forget_partition(PartCtx * c)
{
  std::lock_guard<std::mutex> lock(g_key_mutex);
  for (auto & kv : c->sumtables) g_key_owner.erase(kv.first);
}

static double * sumtable_shadow(PartCtx * c, const double * key, bool create)
{
  auto it = c->sumtables.find(key);
  if (it != c->sumtables.end())
  {
    c->sumtable_stamp[key] = ++c->sumtable_clock;
    return it->second;
  }
  if (!create)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID,
                      "pll_compute_likelihood_derivatives: sumtable was never built.");
    return nullptr;
  }
  /* a caller that frees its sumtable buffers with free() never tells us: cap the shadows, oldest use first */
  while (c->sumtables.size() >= PLLB200_MAX_SUMTABLES)
  {
    const void * victim = nullptr;
    unsigned long long oldest = ~0ull;
    for (auto & kv : c->sumtables)
    {
      const unsigned long long st = c->sumtable_stamp[kv.first];
      if (st < oldest)
      {
        oldest = st;
        victim = kv.first;
      }
    }
    sumtable_forget_key(victim);
    if (c->sumtables.count(victim)) break; /* not ours to drop (cannot happen): do not spin */
  }
  double * d = nullptr;
  if (!pllb200_cuda_ok(cudaMalloc(&d, sizeof(double) * c->clv_stride), "sumtable alloc")) return nullptr;
  {
    std::lock_guard<std::mutex> lock(g_key_mutex);
    auto own = g_key_owner.find(key);
    if (own != g_key_owner.end() && own->second != c)
    {
      /* the same host buffer was used with another partition before: drop that shadow */
      PartCtx * o = own->second;
      auto sh = o->sumtables.find(key);
      if (sh != o->sumtables.end())
      {
        cudaSetDevice(o->device); /* the other partition may live on another GPU */
        cudaStreamSynchronize(o->stream);
        cudaFree(sh->second);
        o->sumtables.erase(sh);
        o->sumtable_stamp.erase(key);
        cudaSetDevice(c->device);
      }
    }
    g_key_owner[key] = c;
  }
  c->sumtables[key] = d;
  c->sumtable_stamp[key] = ++c->sumtable_clock;
  return d;
}

extern "C" int pll_update_sumtable(pll_partition_t * partition, unsigned int parent_clv_index,
                                   unsigned int child_clv_index, int parent_scaler_index,
                                   int child_scaler_index, const unsigned int * params_indices,
                                   double * sumtable)
{
  if (!enter(partition)) return PLL_FAILURE;
  PartCtx * c = ctx_of(partition);
  if (!sumtable)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "NULL sumtable buffer.");
    return PLL_FAILURE;
  }
  double * d = sumtable_shadow(c, sumtable, true);
  if (!d) return PLL_FAILURE;
  return launch_sumtable(c, parent_clv_index, child_clv_index, parent_scaler_index,
                         child_scaler_index, params_indices, d) ? PLL_SUCCESS : PLL_FAILURE;
}

extern "C" int pllb200_newton_branch(pll_partition_t * partition, const unsigned int * params_indices,
                                     const double * sumtable, double xmin, double xguess, double xmax,
                                     double tolerance, unsigned int max_iters, double brlen_scaler,
                                     double * x_out, int * iterations)
{
  /* returns 1 converged, 0 iteration limit (pll_errno = NEWTON_LIMIT), -1 error, -2 not supported for
   * this partition (caller falls back to one derivative call per iteration) */
  if (!enter(partition)) return -1;
  PartCtx * c = ctx_of(partition);
  if (!newton_fused_supported(c)) return -2;
  double * d = sumtable_shadow(c, sumtable, false);
  if (!d) return -1;
  const int slot = dev_alloc_slot(c);
  if (!launch_newton(c, params_indices, d, xmin, xguess, xmax, tolerance, max_iters, brlen_scaler, slot))
    return -1;
  if (!dev_wait_results(c) || !dev_wait_slot(c, slot)) return -1;
  g_d2h_bytes += 4 * sizeof(double);
  const double * r = c->h_result + slot * 4;
  *x_out = r[0];
  if (iterations) *iterations = (int)r[2];
  g_newton_iterations += (unsigned long long)r[2];
  if (r[1] == 0.0) return 1;
  if (r[1] == 1.0)
  {
    pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_LIMIT, "Exceeded maximum number of iterations");
    return 0;
  }
  pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_DERIV, "Wrong likelihood derivatives");
  return -1;
}

/* One Newton iteration loop on a branch length shared by several partitions (linked or scaled branch
 * lengths; replaces the per-iteration derivative callback of pll_optimize.c:1223-1287 + 1395-1560 for that
 * case).  Null partitions are skipped.  Returns like pllb200_newton_branch; -2 if any partition is not
 * eligible for the fused path. */
extern "C" int pllb200_newton_branch_multi(pll_partition_t ** partitions, unsigned int partition_count,
                                           unsigned int ** params_indices, double ** sumtables,
                                           const double * brlen_scalers, double xmin, double xguess,
                                           double xmax, double tolerance, unsigned int max_iters,
                                           double * x_out, int * iterations)
{
  std::vector<PartCtx *> cs;
  std::vector<unsigned int *> pi;
  std::vector<double *> st;
  std::vector<double> sc;
  for (unsigned int p = 0; p < partition_count; ++p)
  {
    if (!partitions[p]) continue;
    if (!enter(partitions[p])) return -1;
    PartCtx * c = ctx_of(partitions[p]);
    if (!newton_fused_supported(c)) return -2; /* an exchange, if any, is the FIRST partition's (site-sharded ranks) */
    double * d = sumtable_shadow(c, sumtables[p], false);
    if (!d) return -1;
    cs.push_back(c);
    pi.push_back(params_indices[p]);
    st.push_back(d);
    sc.push_back(brlen_scalers ? brlen_scalers[p] : 1.0);
  }
  if (cs.empty()) return -2;
  PartCtx * c0 = cs[0];
  const int slot = dev_alloc_slot(c0);
  if (!launch_newton_multi(cs.data(), (unsigned int)cs.size(), pi.data(), st.data(), sc.data(), xmin, xguess,
                           xmax, tolerance, max_iters, slot))
    return -1;
  if (!dev_wait_results(c0) || !dev_wait_slot(c0, slot)) return -1;
  g_d2h_bytes += 4 * sizeof(double);
  const double * r = c0->h_result + slot * 4;
  *x_out = r[0];
  if (iterations) *iterations = (int)r[2];
  g_newton_iterations += (unsigned long long)r[2];
  if (r[1] == 0.0) return 1;
  if (r[1] == 1.0)
  {
    pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_LIMIT, "Exceeded maximum number of iterations");
    return 0;
  }
  pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_DERIV, "Wrong likelihood derivatives");
  return -1;
}

/* Queued variant for a whole smoothing pass: Newton on one branch, then the P-matrix of that branch from
 * the length the kernel left on the device — no host wait.  The start value is `host_x0` on the first
 * visit of a branch and the device value afterwards (`use_device_x0`).  Results are fetched with
 * pllb200_newton_collect.  Returns 1 queued, -1 error, -2 not supported for this partition. */
extern "C" int pllb200_newton_branch_queued(pll_partition_t * partition, const unsigned int * params_indices,
                                            const double * sumtable, double xmin, double host_x0, double xmax,
                                            double tolerance, unsigned int max_iters, double brlen_scaler,
                                            unsigned int matrix_index, int use_device_x0, int revert_on_limit)
{
  if (!enter(partition)) return -1;
  PartCtx * c = ctx_of(partition);
  if (!newton_fused_supported(c)) return -2;
  if (matrix_index >= partition->prob_matrices)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid matrix index %u.", matrix_index);
    return -1;
  }
  if (!c->d_brlen)
  {
    const size_t bytes = sizeof(double) * 4 * partition->prob_matrices;
    if (!pllb200_cuda_ok(cudaMalloc(&c->d_brlen, bytes), "branch length buffer")) return -1;
    if (!pllb200_cuda_ok(cudaMemsetAsync(c->d_brlen, 0, bytes, c->stream), "branch length buffer")) return -1;
  }
  double * d = sumtable_shadow(c, sumtable, false);
  if (!d) return -1;
  double * io = c->d_brlen + (size_t)matrix_index * 4;
  const int slot = dev_alloc_slot(c);
  if (!launch_newton(c, params_indices, d, xmin, host_x0, xmax, tolerance, max_iters, brlen_scaler, slot, io,
                     use_device_x0, revert_on_limit))
    return -1;
  return launch_pmatrix_devlen(c, params_indices, matrix_index, io + 1) ? 1 : -1;
}

/* Wait for the queued work and return, per requested matrix index, the final branch length and the
 * Newton status (0 converged, 1 iteration limit, 2 non-finite derivative). */
extern "C" int pllb200_newton_collect(pll_partition_t * partition, const unsigned int * matrix_indices,
                                      unsigned int count, double * lengths, int * status)
{
  if (!enter(partition)) return PLL_FAILURE;
  PartCtx * c = ctx_of(partition);
  if (!c->d_brlen || !count) return dev_flush(c) ? PLL_SUCCESS : PLL_FAILURE;
  std::vector<double> host((size_t)4 * partition->prob_matrices);
  if (count <= 8)
  {
    for (unsigned int i = 0; i < count; ++i)
      if (!pllb200_cuda_ok(cudaMemcpyAsync(host.data() + (size_t)matrix_indices[i] * 4,
                                           c->d_brlen + (size_t)matrix_indices[i] * 4, 4 * sizeof(double),
                                           cudaMemcpyDeviceToHost, c->stream), "branch lengths D2H"))
        return PLL_FAILURE;
    g_d2h_bytes += (size_t)count * 4 * sizeof(double);
  }
  else
  {
    if (!pllb200_cuda_ok(cudaMemcpyAsync(host.data(), c->d_brlen, host.size() * sizeof(double),
                                         cudaMemcpyDeviceToHost, c->stream), "branch lengths D2H"))
      return PLL_FAILURE;
    g_d2h_bytes += host.size() * sizeof(double);
  }
  if (!dev_flush(c)) return PLL_FAILURE;
  for (unsigned int i = 0; i < count; ++i)
  {
    const double * r = host.data() + (size_t)matrix_indices[i] * 4;
    lengths[i] = r[0];
    status[i] = (int)r[2];
    g_newton_iterations += (unsigned long long)r[3];
  }
  return PLL_SUCCESS;
}

extern "C" int pllb200_likelihood_derivatives_async(pll_partition_t * partition,
                                                    int parent_scaler_index, int child_scaler_index,
                                                    double branch_length,
                                                    const unsigned int * params_indices,
                                                    const double * sumtable, double * d_f,
                                                    double * dd_f)
{
  if (!enter(partition)) return PLL_FAILURE;
  PartCtx * c = ctx_of(partition);
  double * d = sumtable_shadow(c, sumtable, false);
  if (!d) return PLL_FAILURE;
  const int slot = dev_alloc_slot(c);
  if (!launch_derivatives(c, parent_scaler_index, child_scaler_index, branch_length,
                          params_indices, d, slot))
    return PLL_FAILURE;
  c->pending.push_back(PendingResult{d_f, slot, 0});
  c->pending.push_back(PendingResult{dd_f, slot, 1});
  return PLL_SUCCESS;
}

extern "C" int pll_compute_likelihood_derivatives(pll_partition_t * partition,
                                                  int parent_scaler_index, int child_scaler_index,
                                                  double branch_length,
                                                  const unsigned int * params_indices,
                                                  const double * sumtable, double * d_f,
                                                  double * dd_f)
{
  if (!enter(partition)) return PLL_FAILURE;
  PartCtx * c = ctx_of(partition);
  double * d = sumtable_shadow(c, sumtable, false);
  if (!d) return PLL_FAILURE;
  const int slot = dev_alloc_slot(c);
  if (!launch_derivatives(c, parent_scaler_index, child_scaler_index, branch_length,
                          params_indices, d, slot))
    return PLL_FAILURE;
  if (!dev_wait_results(c) || !dev_wait_slot(c, slot)) return PLL_FAILURE;
  g_d2h_bytes += 2 * sizeof(double);
  *d_f = c->h_result[slot * 4 + 0];
  *dd_f = c->h_result[slot * 4 + 1];
  return PLL_SUCCESS;
}
