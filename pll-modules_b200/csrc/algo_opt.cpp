/*
 * algo_opt.cpp — model-parameter optimisation over a whole treeinfo (all partitions in lock step),
 * the callers either side of the likelihood hot path.
 *
 *   pllmod_algo_opt_onedim_treeinfo{,_custom}      src/algorithm/pllmod_algorithm.c:729-853   (Brent)
 *   pllmod_algo_opt_brlen_scalers_treeinfo         pllmod_algorithm.c:855-943
 *   pllmod_algo_opt_subst_rates_treeinfo           pllmod_algorithm.c:945-1134                (L-BFGS-B)
 *   pllmod_algo_opt_frequencies_treeinfo           pllmod_algorithm.c:1136-1290
 *   pllmod_algo_opt_alpha_pinv_treeinfo            pllmod_algorithm.c:1292-1430
 *   pllmod_algo_opt_rates_weights_treeinfo         pllmod_algorithm.c:1432-1840
 *   pllmod_algo_opt_brlen_treeinfo                 pllmod_algorithm.c:1842-1870
 *   objective callbacks                            src/algorithm/algo_callback.c:295-701
 *
 * Every objective evaluation is ONE pllmod_treeinfo_compute_loglh(treeinfo, 0): P-matrices, a full
 * traversal and the edge logL of every partition are queued on the partitions' streams and waited
 * for once, so the L-BFGS-B finite-difference points of all partitions are evaluated as one batch of
 * device work per coordinate.  Model arrays are edited in place on the host (as the reference's
 * callers do) and mirrored lazily: eigen_decomp_valid = 0 triggers the re-decomposition + upload.
 */
#include "internal.h"

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#define ALGO_MIN_WEIGHT_RATIO 0.001 /* pllmod_algorithm.h:28-33 */
#define ALGO_MAX_WEIGHT_RATIO 10
#define ALGO_BFGS_FACTR 1e9
#define ALGO_LBFGSB_ERROR 1.0e-4

namespace
{
/* what the lock-step objective needs to know */
struct OptContext
{
  pllmod_treeinfo_t * ti;
  int mask;                 /* PLLMOD_OPT_PARAM_* selecting the participating partitions */
  bool match_all_bits;      /* (flags & mask) == mask instead of (flags & mask) != 0 */
  unsigned int nparts;      /* participating partitions (local + remote) */
  unsigned int params_index;
  const unsigned int * free_count; /* per participating partition */
  const unsigned int * fixed_idx;  /* per participating partition (frequencies / weights) */
  treeinfo_param_set_cb setter;    /* one-dimensional parameters */
};

inline bool takes_part(const OptContext & c, unsigned int p)
{
  const int f = c.ti->params_to_optimize[p] & c.mask;
  return c.match_all_bits ? f == c.mask : f != 0;
}

/* shared tail of every objective: score, per-partition scores, cross-rank convergence flag */
double finish_objective(const OptContext & c, bool evaluate, double * fx, int * converged,
                        double unconverged)
{
  pllmod_treeinfo_t * ti = c.ti;
  double score = -INFINITY;
  if (evaluate) score = -1 * pllmod_treeinfo_compute_loglh(ti, 0);
  if (fx)
  {
    unsigned int j = 0;
    for (unsigned int p = 0; p < ti->partition_count; ++p)
      if (takes_part(c, p)) fx[j++] = -1 * ti->partition_loglh[p];
  }
  if (converged)
  {
    if (ti->parallel_reduce_cb)
      ti->parallel_reduce_cb(ti->parallel_context, &unconverged, 1, PLLMOD_COMMON_REDUCE_SUM);
    converged[c.nparts] = unconverged > 0. ? 0 : 1;
  }
  return score;
}

/* Brent objective: one scalar per participating partition (algo_callback.c:295-352) */
double objective_onedim(void * ctx, double * x, double * fx, int * converged)
{
  const OptContext & c = *(const OptContext *)ctx;
  pllmod_treeinfo_t * ti = c.ti;
  double unconverged = 0.;
  unsigned int j = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!takes_part(c, p)) continue;
    if (ti->partitions[p] && !(converged && converged[j]))
    {
      unconverged = 1.;
      if (x) c.setter(ti, p, &x[j], 1);
    }
    ++j;
  }
  return finish_objective(c, x != nullptr, fx, converged, unconverged);
}

/* ratios x (relative to the entry `fixed`, which is 1) -> normalised vector */
void ratios_to_simplex(const double * x, unsigned int n, unsigned int fixed, double * out)
{
  double total = 1.0;
  for (unsigned int i = 0; i + 1 < n; ++i) total += x[i];
  unsigned int k = 0;
  for (unsigned int i = 0; i < n; ++i)
    if (i != fixed) out[i] = x[k++] / total;
  out[fixed] = 1.0 / total;
}

/* L-BFGS-B objective for every multi-dimensional parameter family (algo_callback.c:354-701) */
double objective_multidim(void * ctx, double ** x, double * fx, int * converged)
{
  const OptContext & c = *(const OptContext *)ctx;
  pllmod_treeinfo_t * ti = c.ti;
  double unconverged = 0.;
  unsigned int part = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!takes_part(c, p)) continue;
    pll_partition_t * pt = ti->partitions[p];
    if (!pt || (converged && converged[part]))
    {
      ++part;
      continue;
    }
    unconverged = 1.;
    if (!x)
    {
      ++part;
      continue;
    }
    const double * v = x[part];
    switch (c.mask)
    {
      case PLLMOD_OPT_PARAM_SUBST_RATES:
      {
        const unsigned int nrates = pt->states * (pt->states - 1) / 2;
        double * rates = pt->subst_params[c.params_index];
        const int * sym = ti->subst_matrix_symmetries[p];
        if (sym)
        {
          unsigned int k = 0;
          for (unsigned int cls = 0; cls <= c.free_count[part]; ++cls)
          {
            const double val = (cls == (unsigned int)sym[nrates - 1]) ? 1.0 : v[k++];
            for (unsigned int j = 0; j < nrates; ++j)
              if ((unsigned int)sym[j] == cls) rates[j] = val;
          }
        }
        else
          memcpy(rates, v, sizeof(double) * (nrates - 1));
        pt->eigen_decomp_valid[c.params_index] = 0;
        break;
      }
      case PLLMOD_OPT_PARAM_FREQUENCIES:
        ratios_to_simplex(v, pt->states, c.fixed_idx[part], pt->frequencies[c.params_index]);
        pt->eigen_decomp_valid[c.params_index] = 0;
        break;
      case PLLMOD_OPT_PARAM_ALPHA | PLLMOD_OPT_PARAM_PINV:
        ti->alphas[p] = v[0];
        if (!pll_compute_gamma_cats(v[0], pt->rate_cats, pt->rates, ti->gamma_mode[p]))
          return PLL_FAILURE;
        for (unsigned int k = 0; k < pt->rate_cats; ++k)
          if (!pll_update_invariant_sites_proportion(pt, ti->param_indices[p][k], v[1]))
            return PLL_FAILURE;
        break;
      case PLLMOD_OPT_PARAM_FREE_RATES:
        memcpy(pt->rates, v, sizeof(double) * pt->rate_cats);
        break;
      case PLLMOD_OPT_PARAM_RATE_WEIGHTS:
        ratios_to_simplex(v, pt->rate_cats, c.fixed_idx[part], pt->rate_weights);
        break;
      default: assert(0);
    }
    ++part;
  }
  return finish_objective(c, x != nullptr, fx, converged, unconverged);
}

unsigned int reduce_max(pllmod_treeinfo_t * ti, unsigned int v)
{
  if (!ti->parallel_reduce_cb) return v;
  double t = (double)v;
  ti->parallel_reduce_cb(ti->parallel_context, &t, 1, PLLMOD_COMMON_REDUCE_MAX);
  return (unsigned int)t;
}

unsigned int count_parts(const pllmod_treeinfo_t * ti, int mask, bool all_bits)
{
  unsigned int n = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    const int f = ti->params_to_optimize[p] & mask;
    if (all_bits ? f == mask : f != 0) ++n;
  }
  return n;
}

/* per-partition parameter vectors with shared bound arrays, in the layout lbfgsb_multi expects */
struct MultiVectors
{
  std::vector<std::vector<double>> store;
  std::vector<double *> x, lo, hi;
  std::vector<int *> kind;
  std::vector<double> lo0, hi0;
  std::vector<int> kind0;
  std::vector<unsigned int> nfree;
  MultiVectors(unsigned int parts, unsigned int width)
      : store(parts), x(parts, nullptr), lo(parts, nullptr), hi(parts, nullptr), kind(parts, nullptr),
        lo0(width ? width : 1), hi0(width ? width : 1),
        kind0(width ? width : 1, PLLMOD_OPT_LBFGSB_BOUND_BOTH), nfree(parts, 0)
  {
    for (unsigned int i = 0; i < parts; ++i)
    {
      lo[i] = lo0.data();
      hi[i] = hi0.data();
      kind[i] = kind0.data();
    }
  }
  double * local(unsigned int part, unsigned int n)
  {
    store[part].assign(n ? n : 1, 0.0);
    x[part] = store[part].data();
    return x[part];
  }
};

/* one-dimensional getters / setters (pllmod_algorithm.c:554-647) */
int get_alpha(const pllmod_treeinfo_t * ti, unsigned int p, double * v, unsigned int)
{
  if (p >= ti->partition_count) return PLL_FAILURE;
  v[0] = ti->alphas[p];
  return PLL_SUCCESS;
}
int set_alpha(pllmod_treeinfo_t * ti, unsigned int p, const double * v, unsigned int)
{
  if (p >= ti->partition_count) return PLL_FAILURE;
  ti->alphas[p] = v[0];
  pll_partition_t * pt = ti->partitions[p];
  return pll_compute_gamma_cats(v[0], pt->rate_cats, pt->rates, ti->gamma_mode[p]);
}
int get_pinv(const pllmod_treeinfo_t * ti, unsigned int p, double * v, unsigned int)
{
  if (p >= ti->partition_count) return PLL_FAILURE;
  v[0] = ti->partitions[p]->prop_invar[ti->param_indices[p][0]];
  return PLL_SUCCESS;
}
int set_pinv(pllmod_treeinfo_t * ti, unsigned int p, const double * v, unsigned int)
{
  if (p >= ti->partition_count) return PLL_FAILURE;
  pll_partition_t * pt = ti->partitions[p];
  for (unsigned int k = 0; k < pt->rate_cats; ++k)
    if (!pll_update_invariant_sites_proportion(pt, ti->param_indices[p][k], v[0])) return PLL_FAILURE;
  return PLL_SUCCESS;
}
int get_brlen_scaler(const pllmod_treeinfo_t * ti, unsigned int p, double * v, unsigned int)
{
  if (p >= ti->partition_count) return PLL_FAILURE;
  v[0] = ti->brlen_scalers[p];
  return PLL_SUCCESS;
}
int set_brlen_scaler(pllmod_treeinfo_t * ti, unsigned int p, const double * v, unsigned int)
{
  if (p >= ti->partition_count) return PLL_FAILURE;
  ti->brlen_scalers[p] = v[0];
  return PLL_SUCCESS;
}

bool clamp_all_branches(pllmod_treeinfo_t * ti, double blmin, double blmax)
{
  bool changed = false;
  for (unsigned int i = 0; i < ti->subnode_count; ++i)
  {
    pll_unode_t * n = ti->subnodes[i];
    if (n->length < blmin || n->length > blmax)
    {
      pllmod_treeinfo_set_branch_length(ti, n, n->length < blmin ? blmin : blmax);
      changed = true;
    }
  }
  return changed;
}
}  // namespace

/* ---- one-dimensional parameters (Brent, all partitions in lock step) --------------------------- */

extern "C" double pllmod_algo_opt_onedim_treeinfo_custom(pllmod_treeinfo_t * ti, int param_to_optimize,
                                                         treeinfo_param_get_cb getter,
                                                         treeinfo_param_set_cb setter,
                                                         double min_value, double max_value,
                                                         double tolerance)
{
  const unsigned int n = count_parts(ti, param_to_optimize, false);
  if (n > 0)
  {
    std::vector<double> vals(n, 0.0);
    std::vector<int> mask(n, 0);
    unsigned int j = 0;
    for (unsigned int p = 0; p < ti->partition_count; ++p)
    {
      if (!(ti->params_to_optimize[p] & param_to_optimize)) continue;
      if (ti->partitions[p])
      {
        getter(ti, p, &vals[j], 1);
        mask[j] = 1;
      }
      ++j;
    }
    OptContext c{ti, param_to_optimize, false, n, 0, nullptr, nullptr, setter};
    if (!pllmod_opt_minimize_brent_multi(n, mask.data(), &min_value, vals.data(), &max_value,
                                         tolerance, vals.data(), nullptr, nullptr, &c,
                                         objective_onedim, 1))
      return -INFINITY;
  }
  return -1 * pllmod_treeinfo_compute_loglh(ti, 0);
}

extern "C" double pllmod_algo_opt_onedim_treeinfo(pllmod_treeinfo_t * ti, int param_to_optimize,
                                                  double min_value, double max_value,
                                                  double tolerance)
{
  if (__builtin_popcount(param_to_optimize) > 1)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID,
                     "Multi-parameter optimization is not supported by the "
                     "pllmod_algo_opt_onedim_treeinfo() function!");
    return -INFINITY;
  }
  treeinfo_param_get_cb getter;
  treeinfo_param_set_cb setter;
  switch (param_to_optimize)
  {
    case PLLMOD_OPT_PARAM_ALPHA: getter = get_alpha, setter = set_alpha; break;
    case PLLMOD_OPT_PARAM_PINV: getter = get_pinv, setter = set_pinv; break;
    case PLLMOD_OPT_PARAM_BRANCH_LEN_SCALER: getter = get_brlen_scaler, setter = set_brlen_scaler; break;
    default:
      pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Unsupported parameter: %d", param_to_optimize);
      return -INFINITY;
  }
  return pllmod_algo_opt_onedim_treeinfo_custom(ti, param_to_optimize, getter, setter, min_value,
                                                max_value, tolerance);
}

extern "C" double pllmod_algo_opt_brlen_scalers_treeinfo(pllmod_treeinfo_t * ti, double min_scaler,
                                                         double max_scaler, double min_brlen,
                                                         double max_brlen, double lh_epsilon)
{
  if (ti->brlen_linkage != PLLMOD_COMMON_BRLEN_SCALED)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID,
                     "Branch length scaler optimization works only in scaled branch length mode.");
    return (double)PLL_FAILURE;
  }
  const double old_loglh = pllmod_treeinfo_compute_loglh(ti, 0);
  std::vector<double> old_scalers(ti->brlen_scalers, ti->brlen_scalers + ti->partition_count);
  std::vector<double> old_brlen(ti->branch_lengths[0], ti->branch_lengths[0] + ti->tree->edge_count);

  /* bring every scaler into [min, max] by one global factor, compensated on the branches */
  if (ti->parallel_reduce_cb)
  {
    for (unsigned int p = 0; p < ti->partition_count; ++p)
      if (!ti->partitions[p]) ti->brlen_scalers[p] = 0.0;
    ti->parallel_reduce_cb(ti->parallel_context, ti->brlen_scalers, ti->partition_count,
                           PLLMOD_COMMON_REDUCE_MAX);
  }
  double lowest = ti->brlen_scalers[0], highest = ti->brlen_scalers[0];
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    lowest = fmin(lowest, ti->brlen_scalers[p]);
    highest = fmax(highest, ti->brlen_scalers[p]);
  }
  if (lowest < min_scaler || highest > max_scaler)
  {
    const double g = lowest < min_scaler ? min_scaler / lowest : max_scaler / highest;
    for (unsigned int p = 0; p < ti->partition_count; ++p) ti->brlen_scalers[p] *= g;
    pllmod_treeinfo_scale_branches_all(ti, 1.0 / g);
  }

  double loglh = -1 * pllmod_algo_opt_onedim_treeinfo(ti, PLLMOD_OPT_PARAM_BRANCH_LEN_SCALER,
                                                      min_scaler, max_scaler, lh_epsilon);
  pllmod_treeinfo_normalize_brlen_scalers(ti);
  if (clamp_all_branches(ti, min_brlen, max_brlen))
  {
    loglh = pllmod_treeinfo_compute_loglh(ti, 0);
    if (loglh < old_loglh)
    {
      /* worse than before once the branches are legal again: give everything back */
      for (unsigned int p = 0; p < ti->partition_count; ++p)
        if (ti->partitions[p]) ti->brlen_scalers[p] = old_scalers[p];
      for (unsigned int i = 0; i < ti->subnode_count; ++i)
      {
        pll_unode_t * n = ti->subnodes[i];
        if (n->node_index < n->back->node_index)
          pllmod_treeinfo_set_branch_length(ti, n, old_brlen[n->pmatrix_index]);
      }
      loglh = pllmod_treeinfo_compute_loglh(ti, 0);
    }
  }
  return -1 * loglh;
}

/* ---- multi-dimensional parameters (L-BFGS-B, all partitions in lock step) ---------------------- */

extern "C" double pllmod_algo_opt_subst_rates_treeinfo(pllmod_treeinfo_t * ti,
                                                       unsigned int params_index, double min_rate,
                                                       double max_rate, double bfgs_factor,
                                                       double tolerance)
{
  const double factor = bfgs_factor > 0. ? bfgs_factor : ALGO_BFGS_FACTR;
  const unsigned int parts = count_parts(ti, PLLMOD_OPT_PARAM_SUBST_RATES, false);
  if (!parts) return -1 * pllmod_treeinfo_compute_loglh(ti, 0);

  /* free rate classes per partition (the class of the last rate is fixed to 1) */
  std::vector<unsigned int> nfree(parts, 0);
  unsigned int width = 0, part = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!(ti->params_to_optimize[p] & PLLMOD_OPT_PARAM_SUBST_RATES)) continue;
    if (ti->partitions[p])
    {
      const unsigned int nrates = ti->partitions[p]->states * (ti->partitions[p]->states - 1) / 2;
      const int * sym = ti->subst_matrix_symmetries[p];
      unsigned int k = nrates - 1;
      if (sym)
      {
        k = 0;
        for (unsigned int j = 0; j < nrates; ++j)
          if ((unsigned int)sym[j] > k) ++k;
      }
      nfree[part] = k;
      width = std::max(width, k);
    }
    ++part;
  }
  width = reduce_max(ti, width);

  MultiVectors mv(parts, width);
  for (unsigned int k = 0; k < width; ++k) mv.lo0[k] = min_rate, mv.hi0[k] = max_rate;
  part = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!(ti->params_to_optimize[p] & PLLMOD_OPT_PARAM_SUBST_RATES)) continue;
    pll_partition_t * pt = ti->partitions[p];
    if (pt)
    {
      const unsigned int nrates = pt->states * (pt->states - 1) / 2;
      const double * rates = pt->subst_params[params_index];
      const int * sym = ti->subst_matrix_symmetries[p];
      double * x = mv.local(part, nfree[part]);
      unsigned int cls = 0;
      for (unsigned int k = 0; k < nfree[part]; ++k)
      {
        if (sym)
        {
          if ((unsigned int)sym[nrates - 1] == cls) ++cls;
          for (unsigned int j = 0; j < nrates; ++j)
            if ((unsigned int)sym[j] == cls)
            {
              x[k] = rates[j];
              break;
            }
          ++cls;
        }
        else
          x[k] = rates[k];
        x[k] = fmin(fmax(x[k], min_rate), max_rate);
      }
      mv.nfree[part] = nfree[part];
    }
    ++part;
  }
  OptContext c{ti, PLLMOD_OPT_PARAM_SUBST_RATES, false, parts, params_index, mv.nfree.data(), nullptr,
               nullptr};
  return pllmod_opt_minimize_lbfgsb_multi(parts, mv.x.data(), mv.lo.data(), mv.hi.data(),
                                          mv.kind.data(), mv.nfree.data(), width, factor, tolerance,
                                          &c, objective_multidim);
}

extern "C" double pllmod_algo_opt_frequencies_treeinfo(pllmod_treeinfo_t * ti,
                                                       unsigned int params_index, double min_freq,
                                                       double max_freq, double bfgs_factor,
                                                       double tolerance)
{
  const double factor = bfgs_factor > 0. ? bfgs_factor : ALGO_BFGS_FACTR;
  const unsigned int parts = count_parts(ti, PLLMOD_OPT_PARAM_FREQUENCIES, false);
  if (!parts) return -1 * pllmod_treeinfo_compute_loglh(ti, 0);
  unsigned int width = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
    if ((ti->params_to_optimize[p] & PLLMOD_OPT_PARAM_FREQUENCIES) && ti->partitions[p])
      width = std::max(width, ti->partitions[p]->states - 1);
  width = reduce_max(ti, width);

  MultiVectors mv(parts, width);
  for (unsigned int k = 0; k < width; ++k) mv.lo0[k] = min_freq, mv.hi0[k] = max_freq;
  std::vector<unsigned int> fixed(parts, 0);
  unsigned int part = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!(ti->params_to_optimize[p] & PLLMOD_OPT_PARAM_FREQUENCIES)) continue;
    pll_partition_t * pt = ti->partitions[p];
    if (pt)
    {
      const double * fr = pt->frequencies[params_index];
      const unsigned int s = pt->states;
      /* the first state above the lower bound is the reference state of the ratios */
      unsigned int ref_state = s;
      for (unsigned int j = 0; j < s; ++j)
        if (fr[j] > min_freq)
        {
          ref_state = j;
          break;
        }
      assert(ref_state < s);
      double * x = mv.local(part, s - 1);
      unsigned int k = 0;
      for (unsigned int j = 0; j < s; ++j)
        if (j != ref_state) x[k++] = fr[j] / fr[ref_state];
      mv.nfree[part] = s - 1;
      fixed[part] = ref_state;
    }
    ++part;
  }
  OptContext c{ti, PLLMOD_OPT_PARAM_FREQUENCIES, false, parts, params_index, mv.nfree.data(),
               fixed.data(), nullptr};
  return pllmod_opt_minimize_lbfgsb_multi(parts, mv.x.data(), mv.lo.data(), mv.hi.data(),
                                          mv.kind.data(), mv.nfree.data(), width, factor, tolerance,
                                          &c, objective_multidim);
}

extern "C" double pllmod_algo_opt_alpha_pinv_treeinfo(pllmod_treeinfo_t * ti,
                                                      unsigned int params_index, double min_alpha,
                                                      double max_alpha, double min_pinv,
                                                      double max_pinv, double bfgs_factor,
                                                      double tolerance)
{
  const double factor = bfgs_factor > 0. ? bfgs_factor : ALGO_BFGS_FACTR;
  const int mask = PLLMOD_OPT_PARAM_ALPHA | PLLMOD_OPT_PARAM_PINV;
  const unsigned int parts = count_parts(ti, mask, true);
  if (!parts) return -1 * pllmod_treeinfo_compute_loglh(ti, 0);
  MultiVectors mv(parts, 2);
  mv.lo0[0] = min_alpha ? min_alpha : PLLMOD_OPT_MIN_ALPHA;
  mv.hi0[0] = max_alpha ? max_alpha : PLLMOD_OPT_MAX_ALPHA;
  mv.lo0[1] = min_pinv > ALGO_LBFGSB_ERROR ? min_pinv : PLLMOD_OPT_MIN_PINV + ALGO_LBFGSB_ERROR;
  mv.hi0[1] = max_pinv ? max_pinv : PLLMOD_OPT_MAX_PINV;
  unsigned int part = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if ((ti->params_to_optimize[p] & mask) != mask) continue;
    if (ti->partitions[p])
    {
      double * x = mv.local(part, 2);
      x[0] = ti->alphas[p];
      x[1] = ti->partitions[p]->prop_invar[params_index];
      mv.nfree[part] = 2;
    }
    ++part;
  }
  OptContext c{ti, mask, true, parts, params_index, mv.nfree.data(), nullptr, nullptr};
  return pllmod_opt_minimize_lbfgsb_multi(parts, mv.x.data(), mv.lo.data(), mv.hi.data(),
                                          mv.kind.data(), mv.nfree.data(), 2, factor, tolerance, &c,
                                          objective_multidim);
}

extern "C" double pllmod_algo_opt_brlen_treeinfo(pllmod_treeinfo_t * ti, double min_brlen,
                                                 double max_brlen, double lh_epsilon, int max_iters,
                                                 int opt_method, int radius)
{
  return pllmod_opt_optimize_branch_lengths_local_multi(
      ti->partitions, ti->partition_count, ti->root, ti->param_indices, ti->deriv_precomp,
      ti->branch_lengths, ti->brlen_scalers, min_brlen, max_brlen, lh_epsilon, max_iters, radius, 1,
      opt_method, ti->brlen_linkage, ti->parallel_context, ti->parallel_reduce_cb);
}

/* ---- free rates + weights (two-step L-BFGS-B, pllmod_algorithm.c:1432-1840) --------------------- */
namespace
{
const int RW_MASK = PLLMOD_OPT_PARAM_FREE_RATES | PLLMOD_OPT_PARAM_RATE_WEIGHTS;

/* multiply the rates of one partition by f and compensate on its branches so that logL is unchanged */
void scale_rates_and_branches(pllmod_treeinfo_t * ti, unsigned int p, double f)
{
  pll_partition_t * pt = ti->partitions[p];
  for (unsigned int k = 0; k < pt->rate_cats; ++k) pt->rates[k] *= f;
  const double g = 1.0 / f;
  if (ti->partition_count == 1)
    pllmod_treeinfo_scale_branches_all(ti, g);
  else if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
    pllmod_treeinfo_scale_branches_partition(ti, p, g);
  else
  {
    assert(ti->brlen_scalers);
    ti->brlen_scalers[p] *= g;
  }
}

/* enforce sum_k w_k r_k = 1 */
void renormalize_free_rates(pllmod_treeinfo_t * ti)
{
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    pll_partition_t * pt = ti->partitions[p];
    if (!pt || !(ti->params_to_optimize[p] & RW_MASK)) continue;
    double mean = 0.0;
    for (unsigned int k = 0; k < pt->rate_cats; ++k) mean += pt->rates[k] * pt->rate_weights[k];
    scale_rates_and_branches(ti, p, 1.0 / mean);
  }
}
}  // namespace

extern "C" double pllmod_algo_opt_rates_weights_treeinfo(pllmod_treeinfo_t * ti, double min_rate,
                                                         double max_rate, double min_brlen,
                                                         double max_brlen, double bfgs_factor,
                                                         double tolerance)
{
  const double factor = bfgs_factor > 0. ? bfgs_factor : ALGO_BFGS_FACTR;
  const unsigned int parts = count_parts(ti, RW_MASK, false);
  unsigned int width = 0;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
    if ((ti->params_to_optimize[p] & RW_MASK) && ti->partitions[p])
      width = std::max(width, ti->partitions[p]->rate_cats);
  const double old_logl = pllmod_treeinfo_compute_loglh(ti, 0);
  if (!parts) return -1 * old_logl;
  width = reduce_max(ti, width);

  /* state to return to if legalising the branches afterwards makes things worse */
  const bool linked = ti->brlen_linkage != PLLMOD_COMMON_BRLEN_UNLINKED;
  std::vector<std::vector<double>> old_rates(ti->partition_count), old_weights(ti->partition_count);
  std::vector<double> old_brlens, old_scalers;
  if (linked)
  {
    old_brlens.assign(ti->linked_branch_lengths, ti->linked_branch_lengths + ti->tree->edge_count);
    if (ti->brlen_scalers) old_scalers.assign(ti->brlen_scalers, ti->brlen_scalers + ti->partition_count);
  }
  MultiVectors mv(parts, width);
  std::vector<unsigned int> fixed(parts, 0);
  {
    unsigned int part = 0;
    for (unsigned int p = 0; p < ti->partition_count; ++p)
    {
      if (!(ti->params_to_optimize[p] & RW_MASK)) continue;
      pll_partition_t * pt = ti->partitions[p];
      if (pt)
      {
        mv.local(part, width);
        if (linked)
        {
          old_rates[p].assign(pt->rates, pt->rates + pt->rate_cats);
          old_weights[p].assign(pt->rate_weights, pt->rate_weights + pt->rate_cats);
        }
      }
      ++part;
    }
  }

  /* rates outside [min_rate, max_rate]: rescale the whole rate vector (and the branches) */
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    pll_partition_t * pt = ti->partitions[p];
    if (!pt || !(ti->params_to_optimize[p] & PLLMOD_OPT_PARAM_FREE_RATES)) continue;
    double lowest = pt->rates[0], highest = pt->rates[0];
    for (unsigned int k = 1; k < pt->rate_cats; ++k)
    {
      lowest = fmin(lowest, pt->rates[k]);
      highest = fmax(highest, pt->rates[k]);
    }
    if (lowest < min_rate)
      scale_rates_and_branches(ti, p, min_rate / lowest);
    else if (highest > max_rate)
      scale_rates_and_branches(ti, p, max_rate / highest);
  }

  OptContext c{ti, 0, false, parts, 0, mv.nfree.data(), fixed.data(), nullptr};
  double cur_logl = -1 * pllmod_treeinfo_compute_loglh(ti, 0), prev_logl;
  do
  {
    prev_logl = cur_logl;
    /* step 1: weights as ratios to the reference category */
    unsigned int part = 0;
    for (unsigned int p = 0; p < ti->partition_count; ++p)
    {
      if (!(ti->params_to_optimize[p] & PLLMOD_OPT_PARAM_RATE_WEIGHTS)) continue;
      pll_partition_t * pt = ti->partitions[p];
      if (pt)
      {
        const unsigned int K = pt->rate_cats, ref_cat = fixed[part];
        mv.nfree[part] = K - 1;
        unsigned int k = 0;
        for (unsigned int i = 0; i < K; ++i)
        {
          if (i == ref_cat) continue;
          mv.kind0[k] = PLLMOD_OPT_LBFGSB_BOUND_BOTH;
          mv.lo0[k] = ALGO_MIN_WEIGHT_RATIO;
          mv.hi0[k] = ALGO_MAX_WEIGHT_RATIO;
          const double r = pt->rate_weights[i] / pt->rate_weights[ref_cat];
          mv.x[part][k] = fmin(fmax(r, mv.lo0[k]), mv.hi0[k]);
          ++k;
        }
      }
      ++part;
    }
    c.mask = PLLMOD_OPT_PARAM_RATE_WEIGHTS;
    cur_logl = pllmod_opt_minimize_lbfgsb_multi(parts, mv.x.data(), mv.lo.data(), mv.hi.data(),
                                                mv.kind.data(), mv.nfree.data(), width, factor,
                                                tolerance, &c, objective_multidim);
    /* step 2: the rates themselves */
    part = 0;
    for (unsigned int p = 0; p < ti->partition_count; ++p)
    {
      if (!(ti->params_to_optimize[p] & PLLMOD_OPT_PARAM_FREE_RATES)) continue;
      pll_partition_t * pt = ti->partitions[p];
      if (pt)
      {
        const unsigned int K = pt->rate_cats;
        mv.nfree[part] = K;
        for (unsigned int i = 0; i < K; ++i)
        {
          mv.kind0[i] = PLLMOD_OPT_LBFGSB_BOUND_BOTH;
          mv.lo0[i] = min_rate;
          mv.hi0[i] = max_rate;
          mv.x[part][i] = fmin(fmax(pt->rates[i], min_rate), max_rate);
        }
      }
      ++part;
    }
    c.mask = PLLMOD_OPT_PARAM_FREE_RATES;
    cur_logl = pllmod_opt_minimize_lbfgsb_multi(parts, mv.x.data(), mv.lo.data(), mv.hi.data(),
                                                mv.kind.data(), mv.nfree.data(), width, factor,
                                                tolerance, &c, objective_multidim);
  } while (prev_logl - cur_logl > tolerance);

  renormalize_free_rates(ti);
  cur_logl = pllmod_treeinfo_compute_loglh(ti, 0);
  if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_SCALED && ti->partition_count > 1)
    pllmod_treeinfo_normalize_brlen_scalers(ti);
  if (linked && clamp_all_branches(ti, min_brlen, max_brlen))
  {
    const double new_logl = pllmod_treeinfo_compute_loglh(ti, 0);
    if (new_logl < old_logl)
    {
      for (unsigned int p = 0; p < ti->partition_count; ++p)
      {
        pll_partition_t * pt = ti->partitions[p];
        if (!pt || !(ti->params_to_optimize[p] & RW_MASK)) continue;
        memcpy(pt->rates, old_rates[p].data(), sizeof(double) * pt->rate_cats);
        memcpy(pt->rate_weights, old_weights[p].data(), sizeof(double) * pt->rate_cats);
      }
      for (unsigned int i = 0; i < ti->subnode_count; ++i)
        ti->subnodes[i]->length = old_brlens[ti->subnodes[i]->pmatrix_index];
      if (ti->brlen_scalers && !old_scalers.empty())
        memcpy(ti->brlen_scalers, old_scalers.data(), sizeof(double) * ti->partition_count);
      renormalize_free_rates(ti);
      cur_logl = pllmod_treeinfo_compute_loglh(ti, 0);
    }
    else
      cur_logl = new_logl;
  }
  return -1 * cur_logl;
}
