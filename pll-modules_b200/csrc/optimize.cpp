/*
 * optimize.cpp — Newton-Raphson branch-length optimisation on top of the device kernels.
 *
 * Host control flow with the reference's tolerances, bracketing and acceptance rules, so that
 * iteration counts and accepted lengths match:
 *   pllmod_opt_minimize_newton_multi / _newton / _newton_old   src/optimize/opt_algorithms.c:86-384
 *   pllmod_opt_optimize_branch_lengths_local / _iterative      src/optimize/pll_optimize.c:778-1144
 *   pllmod_opt_optimize_branch_lengths_local_multi             pll_optimize.c:1395-1951
 *   pllmod_opt_compute_edge_loglikelihood_multi                pll_optimize.c:1182-1221
 * Per optimised branch the device does: 1 sumtable build, one derivative pass per NR iteration
 * (2 doubles back through mapped pinned memory), 1 P-matrix, and 3 single-operation CLV updates
 * when walking on (pll_optimize.c:1659-1708).
 */
#include "internal.h"

#include <cassert>
#include <cmath>
#include <cstdlib>
#include <chrono>
#include <vector>

#define BETTER_LL_THRESHOLD 1e-13 /* pll_optimize.c:40 */

static inline bool needs_lh_check(int method)
{
  return method == PLLMOD_OPT_BLO_NEWTON_OLDSAFE || method == PLLMOD_OPT_BLO_NEWTON_SAFE;
}

/* ---- scalar minimisers ------------------------------------------------------------------------ */

extern "C" int pllmod_opt_minimize_newton_multi(unsigned int xnum, double xmin, double * x,
                                                double xmax, double tolerance,
                                                unsigned int max_iters, int * converged,
                                                void * params,
                                                void (*deriv_func)(void *, double *, double *, double *))
{
  /* opt_algorithms.c:133-262: damped NR on f = d(-lnL)/dx with a shrinking bracket [xl,xh] and a
   * step cap of xmax/max_iters; all xnum variables advance with ONE derivative call per round */
  std::vector<double> xl(xnum, xmin), xh(xnum, xmax), f(xnum, 0.0), df(xnum, 0.0);
  std::vector<int> own_conv;
  if (!converged)
  {
    own_conv.assign(xnum, 0);
    converged = own_conv.data();
  }
  const double dxmax = xmax / max_iters;
  pll_errno = 0;
  for (unsigned int i = 0; i < xnum; ++i) x[i] = fmax(fmin(x[i], xmax), xmin);

  unsigned int iter = 0;
  bool all_converged = false, error = false;
  while (!all_converged && !error)
  {
    if (iter++ > max_iters)
    {
      pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_LIMIT, "Exceeded maximum number of iterations");
      error = true;
      break;
    }
    deriv_func(params, x, f.data(), df.data());
    all_converged = true;
    for (unsigned int i = 0; i < xnum; ++i)
    {
      if (converged[i]) continue;
      if (!std::isfinite(f[i]) || !std::isfinite(df[i]))
      {
        pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_DERIV, "Wrong likelihood derivatives");
        error = true;
        break;
      }
      double dx;
      if (df[i] > 0.0)
      {
        if (fabs(f[i]) < tolerance)
        {
          converged[i] = 1;
          continue;
        }
        if (f[i] < 0.0)
          xl[i] = x[i];
        else
          xh[i] = x[i];
        dx = -f[i] / df[i];
      }
      else
        dx = -f[i] / fabs(df[i]);
      dx = fmax(fmin(dx, dxmax), -dxmax);
      if (x[i] + dx < xl[i]) dx = xl[i] - x[i];
      if (x[i] + dx > xh[i]) dx = xh[i] - x[i];
      if (fabs(dx) < tolerance)
      {
        converged[i] = 1;
        continue;
      }
      x[i] += dx;
      x[i] = fmax(fmin(x[i], xmax), xmin);
      all_converged = all_converged && converged[i];
    }
  }
  return (all_converged && !error) ? PLL_SUCCESS : PLL_FAILURE;
}

namespace
{
struct NewtonWrap
{
  void * params;
  void (*fn)(void *, double, double *, double *);
};
void newton_wrap_cb(void * p, double * x, double * f, double * df)
{
  NewtonWrap * w = (NewtonWrap *)p;
  w->fn(w->params, x[0], f, df);
}
}  // namespace

extern "C" double pllmod_opt_minimize_newton(double xmin, double xguess, double xmax,
                                             double tolerance, unsigned int max_iters,
                                             void * params,
                                             void (*deriv_func)(void *, double, double *, double *))
{
  NewtonWrap w{params, deriv_func};
  double x = xguess;
  return pllmod_opt_minimize_newton_multi(1, xmin, &x, xmax, tolerance, max_iters, nullptr, &w,
                                          newton_wrap_cb)
             ? x
             : (double)PLL_FAILURE;
}

extern "C" double pllmod_opt_minimize_newton_old(double x1, double xguess, double x2,
                                                 double tolerance, unsigned int max_iters,
                                                 void * params,
                                                 void (*deriv_func)(void *, double, double *, double *))
{
  /* opt_algorithms.c:281-384: the legacy safeguarded NR (bisection whenever the function is concave
   * or the Newton step leaves the bracket) */
  double f, df;
  pll_errno = 0;
  double rts = fmin(fmax(xguess, x1), x2), rts_old = 0.0;
  deriv_func(params, rts, &f, &df);
  if (!std::isfinite(f) || !std::isfinite(df))
  {
    pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_DERIV, "wrong likelihood derivatives");
    return (double)-INFINITY;
  }
  if (df >= 0.0 && fabs(f) < tolerance) return rts;
  double xl, xh;
  if (f < 0.0)
  {
    xl = rts;
    xh = x2;
  }
  else
  {
    xh = rts;
    xl = x1;
  }
  double dx = fabs(xh - xl);
  for (unsigned int i = 1; i <= max_iters; ++i)
  {
    rts_old = rts;
    if (df <= 0.0 || (((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0))
    {
      dx = 0.5 * (xh - xl);
      rts = xl + dx;
      if (xl == rts) return rts;
    }
    else
    {
      dx = f / df;
      const double temp = rts;
      rts -= dx;
      if (temp == rts) return rts;
    }
    if (fabs(dx) < tolerance) return rts_old;
    if (i == max_iters) break;
    if (rts < x1) rts = x1;
    deriv_func(params, rts, &f, &df);
    if (!std::isfinite(f) || !std::isfinite(df))
    {
      pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_DERIV, "Wrong likelihood derivatives [it]");
      return (double)-INFINITY;
    }
    if (df > 0.0 && fabs(f) < tolerance) return rts;
    if (f < 0.0)
      xl = rts;
    else
      xh = rts;
  }
  pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_LIMIT, "Exceeded maximum number of iterations");
  return rts_old;
}

/* ---- helpers shared by the single- and multi-partition drivers ---------------------------------- */

extern "C" double pllmod_opt_compute_edge_loglikelihood_multi(
    pll_partition_t ** partitions, size_t partition_count, unsigned int parent_clv_index,
    int parent_scaler_index, unsigned int child_clv_index, int child_scaler_index,
    unsigned int matrix_index, unsigned int ** const params_indices, double * persite_lnl,
    void * parallel_context, void (*parallel_reduce_cb)(void *, double *, size_t, int))
{
  (void)persite_lnl; /* the reference ignores it too (pll_optimize.c:1207-1214) */
  std::vector<double> part(partition_count, 0.0);
  bool ok = true;
  for (size_t p = 0; p < partition_count; ++p)
    if (partitions[p])
      ok &= pllb200_edge_loglikelihood_async(partitions[p], parent_clv_index, parent_scaler_index,
                                             child_clv_index, child_scaler_index, matrix_index,
                                             params_indices[p], &part[p]) == PLL_SUCCESS;
  for (size_t p = 0; p < partition_count; ++p)
    if (partitions[p]) ok &= pllb200_partition_wait_results(partitions[p]) == PLL_SUCCESS;
  if (!ok) return -INFINITY;
  double total = 0.0;
  for (size_t p = 0; p < partition_count; ++p) total += part[p];
  if (parallel_reduce_cb) parallel_reduce_cb(parallel_context, &total, 1, PLLMOD_COMMON_REDUCE_SUM);
  return total;
}

/* PLLB200_TRACE_BLO=1: host wall time per phase of the smoothing passes on stderr */
struct BloTrace
{
  double sum_s = 0, newton_s = 0, pmat_s = 0, reorient_s = 0;
  unsigned long long branches = 0;
  bool on = getenv("PLLB200_TRACE_BLO") != nullptr;
  static double now()
  {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }
};
static thread_local BloTrace g_blo_trace;

/* re-orient one inner CLV: parent <- (right_child->back, left_child->back)  (pll_optimize.c:748-775) */
static void reorient_clv(pll_partition_t ** partitions, size_t partition_count, pll_unode_t * parent,
                         pll_unode_t * right_child, pll_unode_t * left_child)
{
  pll_operation_t op;
  op.parent_clv_index = parent->clv_index;
  op.parent_scaler_index = parent->scaler_index;
  op.child1_clv_index = right_child->back->clv_index;
  op.child1_matrix_index = right_child->back->pmatrix_index;
  op.child1_scaler_index = right_child->back->scaler_index;
  op.child2_clv_index = left_child->back->clv_index;
  op.child2_matrix_index = left_child->back->pmatrix_index;
  op.child2_scaler_index = left_child->back->scaler_index;
  const double t0 = g_blo_trace.on ? BloTrace::now() : 0.0;
  for (size_t p = 0; p < partition_count; ++p)
    if (partitions[p]) pll_update_partials(partitions[p], &op, 1);
  pllb200_batch_flush(); /* batch mode: the loop only collected; one launch per kernel kind */
  if (g_blo_trace.on) g_blo_trace.reorient_s += BloTrace::now() - t0;
}

/* ---- single partition (legacy NR) ------------------------------------------------------------- */

static void single_deriv_cb(void * parameters, double proposal, double * df, double * ddf)
{
  pll_newton_tree_params_t * p = (pll_newton_tree_params_t *)parameters;
  pll_compute_likelihood_derivatives(p->partition, p->tree->scaler_index,
                                     p->tree->back->scaler_index, proposal, p->params_indices,
                                     p->sumtable, df, ddf);
}

static int optimise_branch_rec(pll_newton_tree_params_t * params, int radius, double * loglh,
                               int keep_update)
{
  pll_unode_t * tr_p = params->tree;
  pll_unode_t * tr_q = tr_p->next;
  pll_unode_t * tr_z = tr_q ? tr_q->next : nullptr;
  const double xorig = tr_p->length;

  if (!pll_update_sumtable(params->partition, tr_p->clv_index, tr_p->back->clv_index,
                           tr_p->scaler_index, tr_p->back->scaler_index, params->params_indices,
                           params->sumtable))
    return PLL_FAILURE;
  double xguess = tr_p->length;
  if (xguess < params->branch_length_min || xguess > params->branch_length_max)
    xguess = PLLMOD_OPT_DEFAULT_BRANCH_LEN;
  const double xres = pllmod_opt_minimize_newton_old(params->branch_length_min, xguess,
                                                     params->branch_length_max, params->tolerance,
                                                     params->max_newton_iters, params, single_deriv_cb);
  if (pll_errno) return PLL_FAILURE;
  tr_p->length = tr_p->back->length = xres;

  if (keep_update && fabs(tr_p->length - xorig) > 1e-10)
  {
    pll_update_prob_matrices(params->partition, params->params_indices, &tr_p->pmatrix_index, &xres, 1);
    if (needs_lh_check(params->opt_method))
    {
      const double eval = pll_compute_edge_loglikelihood(
          params->partition, tr_p->clv_index, tr_p->scaler_index, tr_p->back->clv_index,
          tr_p->back->scaler_index, tr_p->pmatrix_index, params->params_indices, nullptr);
      if (eval >= *loglh)
        *loglh = eval;
      else
      {
        tr_p->length = tr_p->back->length = xorig;
        pll_update_prob_matrices(params->partition, params->params_indices, &tr_p->pmatrix_index,
                                 &tr_p->length, 1);
      }
    }
  }

  if (radius && tr_q && tr_z)
  {
    pll_newton_tree_params_t sub = *params;
    reorient_clv(&params->partition, 1, tr_q, tr_p, tr_z);
    sub.tree = tr_q->back;
    if (!optimise_branch_rec(&sub, radius - 1, loglh, keep_update)) return PLL_FAILURE;
    reorient_clv(&params->partition, 1, tr_z, tr_q, tr_p);
    sub.tree = tr_z->back;
    if (!optimise_branch_rec(&sub, radius - 1, loglh, keep_update)) return PLL_FAILURE;
    reorient_clv(&params->partition, 1, tr_p, tr_z, tr_q);
  }
  return PLL_SUCCESS;
}

extern "C" double pllmod_opt_optimize_branch_lengths_local(
    pll_partition_t * partition, pll_unode_t * tree, const unsigned int * params_indices,
    double branch_length_min, double branch_length_max, double tolerance, int smoothings,
    int radius, int keep_update)
{
  pllmod_reset_error();
  if (radius < PLLMOD_OPT_BRLEN_OPTIMIZE_ALL)
  {
    pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_BAD_RADIUS, "Invalid radius for branch length optimization");
    return (double)PLL_FAILURE;
  }
  double loglh = pll_compute_edge_loglikelihood(partition, tree->back->clv_index,
                                                tree->back->scaler_index, tree->clv_index,
                                                tree->scaler_index, tree->pmatrix_index,
                                                params_indices, nullptr);
  pll_newton_tree_params_t params;
  params.partition = partition;
  params.tree = tree;
  params.params_indices = params_indices;
  params.branch_length_min = branch_length_min > 0 ? branch_length_min : PLLMOD_OPT_MIN_BRANCH_LEN;
  params.branch_length_max = branch_length_max > 0 ? branch_length_max : PLLMOD_OPT_MAX_BRANCH_LEN;
  params.tolerance = branch_length_min > 0 ? branch_length_min / 10.0 : PLLMOD_OPT_TOL_BRANCH_LEN;
  params.opt_method = PLLMOD_OPT_BLO_NEWTON_OLDFAST;
  params.max_newton_iters = 30;
  params.sumtable = (double *)pll_aligned_alloc(64, 32); /* key of the device-resident sumtable */
  if (!params.sumtable)
  {
    pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for bl opt variables");
    return (double)PLL_FAILURE;
  }

  unsigned int iters = (unsigned int)smoothings;
  while (iters)
  {
    double new_loglh = loglh;
    params.tree = tree;
    if (!optimise_branch_rec(&params, radius, &new_loglh, keep_update))
    {
      loglh = PLL_FAILURE;
      break;
    }
    if (radius)
    {
      params.tree = tree->back;
      if (!optimise_branch_rec(&params, radius - 1, &new_loglh, keep_update))
      {
        loglh = PLL_FAILURE;
        break;
      }
    }
    new_loglh = pll_compute_edge_loglikelihood(partition, tree->back->clv_index,
                                               tree->back->scaler_index, tree->clv_index,
                                               tree->scaler_index, tree->pmatrix_index,
                                               params_indices, nullptr);
    if (new_loglh - loglh > new_loglh * BETTER_LL_THRESHOLD)
    {
      --iters;
      if (fabs(new_loglh - loglh) < tolerance) iters = 0;
      loglh = new_loglh;
    }
    else
    {
      pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_WORSE_LK,
                       "Local BL opt converged to a worse likelihood score by %f units",
                       new_loglh - loglh);
      loglh = new_loglh;
      break;
    }
  }
  pll_aligned_free(params.sumtable);
  return -1 * loglh;
}

extern "C" double pllmod_opt_optimize_branch_lengths_iterative(
    pll_partition_t * partition, pll_unode_t * tree, const unsigned int * params_indices,
    double branch_length_min, double branch_length_max, double tolerance, int smoothings,
    int keep_update)
{
  return pllmod_opt_optimize_branch_lengths_local(partition, tree, params_indices,
                                                  branch_length_min, branch_length_max, tolerance,
                                                  smoothings, PLLMOD_OPT_BRLEN_OPTIMIZE_ALL,
                                                  keep_update);
}

/* ---- multi partition -------------------------------------------------------------------------- */


static void multi_deriv_cb(void * parameters, double * proposal, double * df, double * ddf)
{
  /* pll_optimize.c:1223-1287.  Derivative kernels of all local partitions are queued first and
   * waited for together (one stream each), then the chain rule for the brlen scalers. */
  pll_newton_tree_params_multi_t * prm = (pll_newton_tree_params_multi_t *)parameters;
  const bool unlinked = prm->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED;
  const unsigned int P = prm->partition_count;
  if (unlinked)
    for (unsigned int p = 0; p < P; ++p) df[p] = ddf[p] = 0;
  else
    *df = *ddf = 0;
  /* queue the derivative pass of every local partition (one stream each), then wait once per stream */
  std::vector<double> pd(2 * (size_t)P, NAN);
  bool ok = true;
  for (unsigned int p = 0; p < P; ++p)
  {
    if (!prm->partitions[p]) continue;
    const double s = prm->brlen_scalers ? prm->brlen_scalers[p] : 1.;
    const double brlen = s * (unlinked ? proposal[p] : proposal[0]);
    ok = pllb200_likelihood_derivatives_async(prm->partitions[p], prm->tree->scaler_index,
                                              prm->tree->back->scaler_index, brlen,
                                              prm->params_indices[p], prm->precomp_buffers[p],
                                              &pd[2 * p], &pd[2 * p + 1]) && ok;
  }
  for (unsigned int p = 0; p < P; ++p)
  {
    if (!prm->partitions[p]) continue;
    ok = pllb200_partition_wait_results(prm->partitions[p]) && ok;
    const double s = prm->brlen_scalers ? prm->brlen_scalers[p] : 1.;
    const double p_df = ok ? pd[2 * p] : NAN, p_ddf = ok ? pd[2 * p + 1] : NAN;
    if (unlinked)
    {
      df[p] = s * p_df;
      ddf[p] = s * s * p_ddf;
    }
    else
    {
      df[0] += s * p_df;
      ddf[0] += s * s * p_ddf;
    }
  }
  if (prm->parallel_reduce_cb)
  {
    if (unlinked)
    {
      prm->parallel_reduce_cb(prm->parallel_context, df, P, PLLMOD_COMMON_REDUCE_SUM);
      prm->parallel_reduce_cb(prm->parallel_context, ddf, P, PLLMOD_COMMON_REDUCE_SUM);
    }
    else
    {
      double d[2] = {*df, *ddf};
      prm->parallel_reduce_cb(prm->parallel_context, d, 2, PLLMOD_COMMON_REDUCE_SUM);
      *df = d[0];
      *ddf = d[1];
    }
  }
}

static void multi_deriv_cb_old(void * parameters, double proposal, double * df, double * ddf)
{
  multi_deriv_cb(parameters, &proposal, df, ddf);
}

/* A smoothing pass without host waits (single local partition, Newton without likelihood check, fused
 * Newton kernel available): branches are only queued; lengths come back once per pass. */
struct QueuedPass
{
  std::vector<pll_unode_t *> nodes;       /* visited branches in order (the root edge appears twice) */
  std::vector<unsigned char> on_device;   /* by pmatrix index: length already left on the device */
};
static thread_local QueuedPass * t_queued = nullptr;

static void refresh_pmatrix_multi(pll_partition_t ** partitions, size_t partition_count,
                                  unsigned int ** params_indices, double ** brlen_buffers,
                                  double * brlen_scalers, pll_unode_t * node)
{
  unsigned int m = node->pmatrix_index;
  for (size_t p = 0; p < partition_count; ++p)
  {
    if (!partitions[p]) continue;
    double l = (brlen_buffers && brlen_buffers[p]) ? brlen_buffers[p][m] : node->length;
    if (brlen_scalers) l *= brlen_scalers[p];
    pll_update_prob_matrices(partitions[p], params_indices[p], &m, &l, 1);
  }
  pllb200_batch_flush();
}

static int optimise_branch_rec_multi(pll_newton_tree_params_multi_t * prm, int radius,
                                     double * loglh, int keep_update)
{
  pll_unode_t * tr_p = prm->tree;
  pll_unode_t * tr_q = tr_p->next;
  pll_unode_t * tr_z = tr_q ? tr_q->next : nullptr;
  const unsigned int m = tr_p->pmatrix_index;
  const bool unlinked = prm->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED;
  unsigned int xnum;
  double xorig_linked, xguess_linked;
  double *xorig, *xguess;

  if (unlinked)
  {
    xnum = prm->partition_count;
    xorig = prm->brlen_orig;
    xguess = prm->brlen_guess;
    for (unsigned int p = 0; p < xnum; ++p)
      xguess[p] = prm->partitions[p] ? prm->brlen_buffers[p][m] : 0.0;
    if (xnum > 1 && prm->parallel_reduce_cb)
      prm->parallel_reduce_cb(prm->parallel_context, xguess, xnum, PLLMOD_COMMON_REDUCE_MAX);
    memcpy(xorig, xguess, sizeof(double) * xnum);
  }
  else
  {
    xnum = 1;
    xorig = &xorig_linked;
    xguess = &xguess_linked;
    *xorig = *xguess = tr_p->length;
  }
  if (prm->converged) memset(prm->converged, 0, sizeof(int) * xnum);

  const double tr0 = g_blo_trace.on ? BloTrace::now() : 0.0;
  for (unsigned int p = 0; p < prm->partition_count; ++p)
    if (prm->partitions[p])
      if (!pll_update_sumtable(prm->partitions[p], tr_p->clv_index, tr_p->back->clv_index,
                               tr_p->scaler_index, tr_p->back->scaler_index,
                               prm->params_indices[p], prm->precomp_buffers[p]))
        return PLL_FAILURE;
  if (!pllb200_batch_flush()) return PLL_FAILURE;
  const double tr1 = g_blo_trace.on ? BloTrace::now() : 0.0;

  const double xmin = prm->branch_length_min, xmax = prm->branch_length_max, xtol = prm->tolerance;
  int retval;
  bool queued = false;
  if (t_queued)
  {
    if (t_queued->on_device.size() <= m) t_queued->on_device.resize(m + 1, 0);
    const double s = prm->brlen_scalers ? prm->brlen_scalers[0] : 1.;
    const int rc = pllb200_newton_branch_queued(prm->partitions[0], prm->params_indices[0],
                                                prm->precomp_buffers[0], xmin, xguess[0], xmax, xtol,
                                                (unsigned int)prm->max_newton_iters, s, m,
                                                t_queued->on_device[m], 1);
    if (rc == -1) return PLL_FAILURE;
    if (rc == 1)
    {
      t_queued->on_device[m] = 1;
      t_queued->nodes.push_back(tr_p);
      queued = true;
    }
    else
      t_queued = nullptr; /* not supported after all: finish the pass the ordinary way */
  }
  if (!queued)
  {
  switch (prm->opt_method)
  {
    case PLLMOD_OPT_BLO_NEWTON_FAST:
    case PLLMOD_OPT_BLO_NEWTON_SAFE:
    {
      /* one local partition and nothing to reduce across ranks: the whole iteration runs on the
       * device (k_newton.cu); otherwise one derivative pass per iteration and partition */
      int fused = -2;
      if (xnum == 1 && prm->partition_count == 1 && prm->partitions[0] &&
          (!prm->parallel_reduce_cb || pllb200_exchange_ranks(prm->partitions[0]) > 1))
      {
        pll_errno = 0;
        const double s = prm->brlen_scalers ? prm->brlen_scalers[0] : 1.;
        fused = pllb200_newton_branch(prm->partitions[0], prm->params_indices[0],
                                      prm->precomp_buffers[0], xmin, xguess[0], xmax, xtol,
                                      (unsigned int)prm->max_newton_iters, s, &xguess[0], nullptr);
        if (fused >= 0 && prm->converged) prm->converged[0] = fused == 1;
      }
      else if (xnum == 1 && !unlinked && prm->partition_count > 1 &&
               (!prm->parallel_reduce_cb || (prm->partitions[0] && pllb200_exchange_ranks(prm->partitions[0]) > 1)))
      {
        /* one shared length over several local partitions: all partitions' derivatives per iteration in
         * one cooperative launch (k_newton.cu, newton_multi_kernel) */
        pll_errno = 0;
        fused = pllb200_newton_branch_multi(prm->partitions, prm->partition_count, prm->params_indices,
                                            prm->precomp_buffers, prm->brlen_scalers, xmin, xguess[0], xmax,
                                            xtol, (unsigned int)prm->max_newton_iters, &xguess[0], nullptr);
        if (fused >= 0 && prm->converged) prm->converged[0] = fused == 1;
      }
      if (fused == -2 && pllb200_batch_active())
        pllb200_batch_end(); /* the derivative kernels of the fallback are launched per partition */
      if (fused == -2)
        retval = pllmod_opt_minimize_newton_multi(xnum, xmin, xguess, xmax, xtol,
                                                  prm->max_newton_iters, prm->converged, prm,
                                                  multi_deriv_cb);
      else
        retval = fused == 1 ? PLL_SUCCESS : PLL_FAILURE;
      break;
    }
    case PLLMOD_OPT_BLO_NEWTON_OLDFAST:
    case PLLMOD_OPT_BLO_NEWTON_OLDSAFE:
      assert(!unlinked);
      xguess[0] = pllmod_opt_minimize_newton_old(xmin, xguess[0], xmax, xtol,
                                                 prm->max_newton_iters, prm, multi_deriv_cb_old);
      retval = pll_errno ? PLL_FAILURE : PLL_SUCCESS;
      break;
    default:
      pllmod_set_error(PLLMOD_ERROR_NOT_IMPLEMENTED, "Optimization method not implemented");
      return PLL_FAILURE;
  }
  if (!retval)
  {
    if (pll_errno != PLLMOD_OPT_ERROR_NEWTON_LIMIT) return PLL_FAILURE;
    /* not converged: keep the last iterate only if a likelihood check will guard it
     * (pll_optimize.c:1524-1549) */
    for (unsigned int p = 0; p < xnum; ++p)
    {
      if (prm->converged && prm->converged[p]) continue;
      if (!needs_lh_check(prm->opt_method)) xguess[p] = xorig[p];
    }
    pllmod_reset_error();
  }

  const double tr2 = g_blo_trace.on ? BloTrace::now() : 0.0;
  bool apply_change = false;
  for (unsigned int p = 0; p < xnum; ++p)
  {
    assert(xguess[p] >= xmin && xguess[p] <= xmax);
    if (fabs(xguess[p] - xorig[p]) < 1e-10) continue;
    apply_change = true;
    if (prm->brlen_buffers && prm->brlen_buffers[p]) prm->brlen_buffers[p][m] = xguess[p];
  }
  if (apply_change)
  {
    if (!unlinked) tr_p->length = tr_p->back->length = xguess[0];
    if (keep_update)
      refresh_pmatrix_multi(prm->partitions, prm->partition_count, prm->params_indices,
                            prm->brlen_buffers, prm->brlen_scalers, tr_p);
    if (needs_lh_check(prm->opt_method))
    {
      assert(keep_update);
      const double eval = pllmod_opt_compute_edge_loglikelihood_multi(
          prm->partitions, prm->partition_count, tr_p->clv_index, tr_p->scaler_index,
          tr_p->back->clv_index, tr_p->back->scaler_index, tr_p->pmatrix_index,
          prm->params_indices, nullptr, prm->parallel_context, prm->parallel_reduce_cb);
      if (eval >= *loglh)
        *loglh = eval;
      else
      {
        for (unsigned int p = 0; p < xnum; ++p)
          if (prm->brlen_buffers && prm->brlen_buffers[p]) prm->brlen_buffers[p][m] = xorig[p];
        if (!unlinked) tr_p->length = tr_p->back->length = xorig[0];
        refresh_pmatrix_multi(prm->partitions, prm->partition_count, prm->params_indices,
                              prm->brlen_buffers, prm->brlen_scalers, tr_p);
      }
    }
  }

  if (g_blo_trace.on)
  {
    g_blo_trace.sum_s += tr1 - tr0;
    g_blo_trace.newton_s += tr2 - tr1;
    g_blo_trace.pmat_s += BloTrace::now() - tr2;
    ++g_blo_trace.branches;
  }
  } /* !queued */

  if (radius && tr_q && tr_z)
  {
    pll_newton_tree_params_multi_t sub = *prm;
    reorient_clv(prm->partitions, prm->partition_count, tr_q, tr_p, tr_z);
    sub.tree = tr_q->back;
    if (!optimise_branch_rec_multi(&sub, radius - 1, loglh, keep_update)) return PLL_FAILURE;
    reorient_clv(prm->partitions, prm->partition_count, tr_z, tr_q, tr_p);
    sub.tree = tr_z->back;
    if (!optimise_branch_rec_multi(&sub, radius - 1, loglh, keep_update)) return PLL_FAILURE;
    reorient_clv(prm->partitions, prm->partition_count, tr_p, tr_z, tr_q);
  }
  return PLL_SUCCESS;
}

static bool collect_queued(QueuedPass & q, pll_partition_t * partition, double ** brlen_buffers)
{
  if (q.nodes.empty()) return true;
  std::vector<unsigned int> idx(q.nodes.size());
  for (size_t i = 0; i < q.nodes.size(); ++i) idx[i] = q.nodes[i]->pmatrix_index;
  std::vector<double> len(idx.size());
  std::vector<int> st(idx.size());
  const int rc = pllb200_newton_collect(partition, idx.data(), (unsigned int)idx.size(), len.data(), st.data());
  bool ok = rc == PLL_SUCCESS;
  for (size_t i = 0; rc == PLL_SUCCESS && i < q.nodes.size(); ++i)
  {
    pll_unode_t * n = q.nodes[i];
    n->length = n->back->length = len[i];
    if (brlen_buffers && brlen_buffers[0]) brlen_buffers[0][idx[i]] = len[i];
    if (st[i] == 2)
    {
      pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_DERIV, "Wrong likelihood derivatives");
      ok = false;
    }
  }
  q.nodes.clear();
  return ok;
}

extern "C" double pllmod_opt_optimize_branch_lengths_local_multi(
    pll_partition_t ** partitions, size_t partition_count, pll_unode_t * tree,
    unsigned int ** params_indices, double ** precomp_buffers, double ** brlen_buffers,
    double * brlen_scalers, double branch_length_min, double branch_length_max,
    double lh_epsilon, int max_iters, int radius, int keep_update, int opt_method,
    int brlen_linkage, void * parallel_context,
    void (*parallel_reduce_cb)(void *, double *, size_t, int))
{
  double result = (double)PLL_FAILURE;
  pllmod_reset_error();
  if (opt_method == PLLMOD_OPT_BLO_NEWTON_FALLBACK || opt_method == PLLMOD_OPT_BLO_NEWTON_GLOBAL)
  {
    pllmod_set_error(PLLMOD_ERROR_NOT_IMPLEMENTED,
                     "Optimization method not implemented: NEWTON_FALLBACK, NEWTON_GLOBAL");
    return (double)PLL_FAILURE;
  }
  if (radius < PLLMOD_OPT_BRLEN_OPTIMIZE_ALL)
  {
    pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_BAD_RADIUS, "Invalid radius for branch length optimization");
    return (double)PLL_FAILURE;
  }
  refresh_pmatrix_multi(partitions, partition_count, params_indices, brlen_buffers, brlen_scalers, tree);
  double loglh = pllmod_opt_compute_edge_loglikelihood_multi(
      partitions, partition_count, tree->back->clv_index, tree->back->scaler_index,
      tree->clv_index, tree->scaler_index, tree->pmatrix_index, params_indices, nullptr,
      parallel_context, parallel_reduce_cb);

  pll_newton_tree_params_multi_t prm;
  memset(&prm, 0, sizeof(prm));
  prm.partitions = partitions;
  prm.partition_count = (unsigned int)partition_count;
  prm.tree = tree;
  prm.params_indices = params_indices;
  prm.branch_length_min = branch_length_min > 0 ? branch_length_min : PLLMOD_OPT_MIN_BRANCH_LEN;
  prm.branch_length_max = branch_length_max > 0 ? branch_length_max : PLLMOD_OPT_MAX_BRANCH_LEN;
  prm.tolerance = branch_length_min > 0 ? branch_length_min / 10.0 : PLLMOD_OPT_TOL_BRANCH_LEN;
  prm.precomp_buffers = precomp_buffers;
  prm.brlen_buffers = brlen_buffers;
  prm.brlen_scalers = brlen_scalers;
  prm.opt_method = opt_method;
  prm.brlen_linkage = partition_count > 1 ? brlen_linkage : PLLMOD_COMMON_BRLEN_LINKED;
  prm.max_newton_iters = 30;
  prm.parallel_context = parallel_context;
  prm.parallel_reduce_cb = parallel_reduce_cb;

  /* buffers the caller did not provide (pll_optimize.c:1324-1393); sumtables are device-side, the
   * host allocation is only the 64-byte key */
  QueuedPass queue;
  std::vector<double *> own_precomp;
  std::vector<int> conv;
  std::vector<double> orig, guess;
  if (!prm.precomp_buffers)
  {
    own_precomp.assign(partition_count, nullptr);
    for (size_t p = 0; p < partition_count; ++p)
      if (partitions[p]) own_precomp[p] = (double *)pll_aligned_alloc(64, 32);
    prm.precomp_buffers = own_precomp.data();
  }
  if (prm.brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
  {
    if (!brlen_buffers)
    {
      pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Unlinked branch lengths need brlen_buffers");
      goto cleanup;
    }
    conv.assign(partition_count, 0);
    orig.assign(partition_count, 0.0);
    guess.assign(partition_count, 0.0);
    prm.converged = conv.data();
    prm.brlen_orig = orig.data();
    prm.brlen_guess = guess.data();
  }

  {
    /* queue whole passes when nothing between two branches needs a number on the host */
    const bool can_queue = keep_update && opt_method == PLLMOD_OPT_BLO_NEWTON_FAST && partition_count == 1 &&
                           partitions[0] && (!parallel_reduce_cb || pllb200_exchange_ranks(partitions[0]) > 1) &&
                           g_opt_fused_newton != 0 &&
                           g_opt_queued_newton != 0;
    unsigned int iters = (unsigned int)max_iters;
    while (iters)
    {
      double new_loglh = loglh;
      prm.tree = tree;
      queue.nodes.clear();
      queue.on_device.assign(queue.on_device.size(), 0);
      t_queued = can_queue ? &queue : nullptr;
      /* many small partitions with one shared length per branch: one launch per kernel kind and branch
       * instead of one per partition (device_ctx.cu, batch mode) */
      if (keep_update && opt_method == PLLMOD_OPT_BLO_NEWTON_FAST && partition_count > 1 &&
          (!parallel_reduce_cb || (partitions[0] && pllb200_exchange_ranks(partitions[0]) > 1)) &&
          prm.brlen_linkage != PLLMOD_COMMON_BRLEN_UNLINKED && g_opt_fused_newton != 0 &&
          pllb200_batch_begin(partitions, (unsigned int)partition_count) < 0)
        goto cleanup;
      int ok = optimise_branch_rec_multi(&prm, radius, &new_loglh, keep_update);
      if (ok && radius)
      {
        prm.tree = tree->back;
        ok = optimise_branch_rec_multi(&prm, radius - 1, &new_loglh, keep_update);
      }
      t_queued = nullptr;
      if (!pllb200_batch_end()) ok = 0;
      /* lengths the device decided on -> tree and caller's buffers (also after a failure: the P-matrices
       * on the device already use them) */
      if (!collect_queued(queue, partitions[0], brlen_buffers) || !ok) goto cleanup;
      new_loglh = pllmod_opt_compute_edge_loglikelihood_multi(
          partitions, partition_count, tree->back->clv_index, tree->back->scaler_index,
          tree->clv_index, tree->scaler_index, tree->pmatrix_index, params_indices, nullptr,
          parallel_context, parallel_reduce_cb);
      if (new_loglh - loglh > new_loglh * BETTER_LL_THRESHOLD)
      {
        --iters;
        if (fabs(new_loglh - loglh) < lh_epsilon) iters = 0;
        loglh = new_loglh;
      }
      else
      {
        if (needs_lh_check(prm.opt_method))
          assert(new_loglh - loglh > new_loglh * BETTER_LL_THRESHOLD);
        pllmod_set_error(PLLMOD_OPT_ERROR_NEWTON_WORSE_LK,
                         "BL opt converged to a worse likelihood score by %.15f units",
                         new_loglh - loglh);
        goto cleanup;
      }
    }
    result = -1 * loglh;
  }
cleanup:
  if (g_blo_trace.on && g_blo_trace.branches)
  {
    fprintf(stderr, "[pllb200 blo] %llu branches: sumtable %.1f us, newton %.1f us, pmatrix %.1f us, reorient %.1f us per branch\n",
            g_blo_trace.branches, 1e6 * g_blo_trace.sum_s / g_blo_trace.branches,
            1e6 * g_blo_trace.newton_s / g_blo_trace.branches, 1e6 * g_blo_trace.pmat_s / g_blo_trace.branches,
            1e6 * g_blo_trace.reorient_s / g_blo_trace.branches);
    g_blo_trace = BloTrace();
  }
  for (double * b : own_precomp) pll_aligned_free(b);
  return result;
}
