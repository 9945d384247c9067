/*
 * opt_params.cpp — single-partition model-parameter optimisation over pll_optimize_options_t
 * (SURVEY.md §8a rows a13/a14): the packed parameter vector <-> partition mapping and the two
 * high-level drivers
 *   pllmod_opt_optimize_onedim     src/optimize/pll_optimize.c:411-452   (Brent)
 *   pllmod_opt_optimize_multidim   pll_optimize.c:473-741                (L-BFGS-B)
 *   pllmod_opt_derivative_func     pll_optimize.c:1146-1159
 * Every objective evaluation = setters on the partition (host) -> P-matrices -> traversal ->
 * edge/root logL on the device (no CLV ever leaves HBM).
 *
 * Packed vector order (pll_optimize.c:72-300): substitution rates (last one fixed to 1, or free
 * classes of the symmetry vector with the class of the last rate fixed) | frequency ratios to the
 * most frequent state | pinv | alpha | free rates | rate-weight ratios | all branches | one branch.
 */
#include "internal.h"

#include <cassert>
#include <cmath>
#include <vector>

namespace
{
#define LBFGSB_ERROR 1.0e-4

unsigned int branch_count(const pll_optimize_options_t * o)
{
  const unsigned int tips = o->lk_params.partition->tips;
  return o->lk_params.rooted ? 2 * tips - 2 : 2 * tips - 3;
}
unsigned int inner_count(const pll_optimize_options_t * o)
{
  const unsigned int tips = o->lk_params.partition->tips;
  return o->lk_params.rooted ? tips - 1 : tips - 2;
}
unsigned int free_rate_classes(const pll_optimize_options_t * o)
{
  const pll_partition_t * part = o->lk_params.partition;
  const int nrates = (int)(part->states * (part->states - 1) / 2);
  if (!o->subst_params_symmetries) return (unsigned int)(nrates - 1);
  int mx = o->subst_params_symmetries[0];
  for (int i = 1; i < nrates; ++i) mx = std::max(mx, o->subst_params_symmetries[i]);
  return (unsigned int)mx;
}

/* push the packed vector into the partition and refresh what depends on it */
int apply_vector(pll_optimize_options_t * o, const double * x)
{
  pll_partition_t * part = o->lk_params.partition;
  const unsigned int which = o->which_parameters;
  const unsigned int * pidx = o->lk_params.params_indices;
  const unsigned int states = part->states, K = part->rate_cats;
  const double * cur = x;

  if (which & PLLMOD_OPT_PARAM_SUBST_RATES)
  {
    const int nrates = (int)(states * (states - 1) / 2);
    std::vector<double> rates(nrates);
    const int * sym = o->subst_params_symmetries;
    if (sym)
    {
      const int classes = (int)free_rate_classes(o);
      int k = 0;
      for (int cls = 0; cls <= classes; ++cls)
      {
        const double val = (cls == sym[nrates - 1]) ? 1.0 : cur[k++];
        for (int j = 0; j < nrates; ++j)
          if (sym[j] == cls) rates[j] = val;
      }
      cur += classes;
    }
    else
    {
      for (int j = 0; j + 1 < nrates; ++j) rates[j] = cur[j];
      rates[nrates - 1] = 1.0;
      cur += nrates - 1;
    }
    pll_set_subst_params(part, o->params_index, rates.data());
  }
  if (which & PLLMOD_OPT_PARAM_FREQUENCIES)
  {
    std::vector<double> freqs(states);
    double total = 1.0;
    for (unsigned int i = 0; i + 1 < states; ++i) total += cur[i];
    unsigned int k = 0;
    for (unsigned int i = 0; i < states; ++i)
      if (i != o->highest_freq_state) freqs[i] = cur[k++] / total;
    freqs[o->highest_freq_state] = 1.0 / total;
    pll_set_frequencies(part, o->params_index, freqs.data());
    cur += states - 1;
  }
  if (which & PLLMOD_OPT_PARAM_PINV)
  {
    for (unsigned int k = 0; k < K; ++k)
      if (!pll_update_invariant_sites_proportion(part, pidx[k], cur[0])) return PLL_FAILURE;
    ++cur;
  }
  if (which & PLLMOD_OPT_PARAM_ALPHA)
  {
    std::vector<double> cats(K);
    o->lk_params.alpha_value = cur[0];
    if (!pll_compute_gamma_cats(cur[0], K, cats.data(), PLL_GAMMA_RATES_MEAN)) return PLL_FAILURE;
    pll_set_category_rates(part, cats.data());
    ++cur;
  }
  if (which & PLLMOD_OPT_PARAM_FREE_RATES)
  {
    pll_set_category_rates(part, cur);
    cur += K;
  }
  if (which & PLLMOD_OPT_PARAM_RATE_WEIGHTS)
  {
    std::vector<double> wts(K);
    double total = 1.0;
    for (unsigned int i = 0; i + 1 < K; ++i) total += cur[i];
    unsigned int k = 0;
    for (unsigned int i = 0; i < K; ++i)
      if (i != o->highest_weight_state) wts[i] = cur[k++] / total;
    wts[o->highest_weight_state] = 1.0 / total;
    pll_set_category_weights(part, wts.data());
    cur += K - 1;
  }
  if (which & PLLMOD_OPT_PARAM_BRANCHES_ALL)
  {
    const unsigned int nb = branch_count(o);
    for (unsigned int i = 0; i < nb; ++i) o->lk_params.branch_lengths[i] = cur[i];
    cur += nb;
  }
  if (which & PLLMOD_OPT_PARAM_BRANCHES_SINGLE)
  {
    /* only the edge under evaluation changes: one P-matrix, no traversal */
    o->lk_params.branch_lengths[0] = cur[0];
    pll_update_prob_matrices(part, pidx, &o->lk_params.where.unrooted_t.edge_pmatrix_index, cur, 1);
    ++cur;
  }
  else
  {
    pll_update_prob_matrices(part, pidx, o->lk_params.matrix_indices, o->lk_params.branch_lengths,
                             branch_count(o));
    pll_update_partials(part, o->lk_params.operations, inner_count(o));
  }
  return PLL_SUCCESS;
}

double negative_loglh(void * p, double * x)
{
  pll_optimize_options_t * o = (pll_optimize_options_t *)p;
  if (x && !apply_vector(o, x)) return -INFINITY;
  pll_partition_t * part = o->lk_params.partition;
  if (o->lk_params.rooted)
    return -pll_compute_root_loglikelihood(part, o->lk_params.where.rooted_t.root_clv_index,
                                           o->lk_params.where.rooted_t.scaler_index,
                                           o->lk_params.params_indices, nullptr);
  return -pll_compute_edge_loglikelihood(part, o->lk_params.where.unrooted_t.parent_clv_index,
                                         o->lk_params.where.unrooted_t.parent_scaler_index,
                                         o->lk_params.where.unrooted_t.child_clv_index,
                                         o->lk_params.where.unrooted_t.child_scaler_index,
                                         o->lk_params.where.unrooted_t.edge_pmatrix_index,
                                         o->lk_params.params_indices, nullptr);
}

double brent_objective(void * p, double x) { return negative_loglh(p, &x); }
}  // namespace

extern "C" void pllmod_opt_derivative_func(void * parameters, double proposal, double * df,
                                           double * ddf)
{
  pll_optimize_options_t * o = (pll_optimize_options_t *)parameters;
  pll_compute_likelihood_derivatives(o->lk_params.partition,
                                     o->lk_params.where.unrooted_t.parent_scaler_index,
                                     o->lk_params.where.unrooted_t.child_scaler_index, proposal,
                                     o->lk_params.params_indices, o->sumtable, df, ddf);
}

extern "C" double pllmod_opt_optimize_onedim(pll_optimize_options_t * o, double umin, double umax)
{
  double lo, hi, guess;
  switch (o->which_parameters)
  {
    case PLLMOD_OPT_PARAM_ALPHA:
      guess = o->lk_params.alpha_value;
      lo = umin > 0 ? umin : PLLMOD_OPT_MIN_ALPHA;
      hi = umax > 0 ? umax : PLLMOD_OPT_MAX_ALPHA;
      break;
    case PLLMOD_OPT_PARAM_PINV:
      guess = o->lk_params.partition->prop_invar[o->params_index];
      lo = umin > 0 ? umin : PLLMOD_OPT_MIN_PINV;
      hi = umax > 0 ? umax : PLLMOD_OPT_MAX_PINV;
      break;
    case PLLMOD_OPT_PARAM_BRANCHES_SINGLE:
      guess = o->lk_params.branch_lengths[0];
      lo = umin > 0 ? umin : PLLMOD_OPT_MIN_BRANCH_LEN;
      hi = umax > 0 ? umax : PLLMOD_OPT_MAX_BRANCH_LEN;
      break;
    default: return -INFINITY; /* not a single scalar parameter */
  }
  double score = 0.0, curvature = 0.0;
  double best = pllmod_opt_minimize_brent(lo, guess, hi, o->pgtol, &score, &curvature, o,
                                          brent_objective);
  apply_vector(o, &best);
  return score;
}

extern "C" double pllmod_opt_optimize_multidim(pll_optimize_options_t * o, double * umin,
                                               double * umax)
{
  pll_partition_t * part = o->lk_params.partition;
  const unsigned int which = o->which_parameters;
  const unsigned int states = part->states, K = part->rate_cats;
  assert(!((which & PLLMOD_OPT_PARAM_BRANCHES_ALL) && (which & PLLMOD_OPT_PARAM_BRANCHES_SINGLE)));
  if (which & PLLMOD_OPT_PARAM_TOPOLOGY)
  {
    pllmod_set_error(PLLMOD_OPT_ERROR_LBFGSB_UNKNOWN, "Topology optimization is not implemented");
    return -INFINITY;
  }

  std::vector<double> x, lo, hi;
  std::vector<int> kind;
  const double * ul = umin;
  const double * uu = umax;
  auto add = [&](double value, double dlo, double dhi, int bound) {
    x.push_back(value);
    lo.push_back(ul ? *ul++ : dlo);
    hi.push_back(uu ? *uu++ : dhi);
    kind.push_back(bound);
  };

  if (which & PLLMOD_OPT_PARAM_SUBST_RATES)
  {
    const unsigned int nrates = states * (states - 1) / 2;
    const unsigned int nfree = free_rate_classes(o);
    const int * sym = o->subst_params_symmetries;
    int cls = 0;
    for (unsigned int i = 0; i < nfree; ++i)
    {
      unsigned int j = i;
      if (sym)
      {
        /* representative rate of the next free class (the class of the last rate is fixed) */
        if (sym[nrates - 1] == cls) ++cls;
        for (j = 0; j < nrates; ++j)
          if (sym[j] == cls) break;
        ++cls;
      }
      add(part->subst_params[o->params_index][j], PLLMOD_OPT_MIN_SUBST_RATE,
          PLLMOD_OPT_MAX_SUBST_RATE, PLLMOD_OPT_LBFGSB_BOUND_BOTH);
    }
  }
  if (which & PLLMOD_OPT_PARAM_FREQUENCIES)
  {
    const double * fr = part->frequencies[o->params_index];
    /* pll_optimize.c:573-576: the scan starts from state 3 */
    o->highest_freq_state = 3;
    for (unsigned int i = 1; i < states; ++i)
      if (fr[i] > fr[o->highest_freq_state]) o->highest_freq_state = i;
    for (unsigned int i = 0; i < states; ++i)
      if (i != o->highest_freq_state)
        add(fr[i] / fr[o->highest_freq_state], PLLMOD_OPT_MIN_FREQ, PLLMOD_OPT_MAX_FREQ,
            PLLMOD_OPT_LBFGSB_BOUND_BOTH);
  }
  if (which & PLLMOD_OPT_PARAM_PINV)
    add(part->prop_invar[o->params_index], PLLMOD_OPT_MIN_PINV + LBFGSB_ERROR, PLLMOD_OPT_MAX_PINV,
        PLLMOD_OPT_LBFGSB_BOUND_BOTH);
  if (which & PLLMOD_OPT_PARAM_ALPHA)
    add(o->lk_params.alpha_value, PLLMOD_OPT_MIN_ALPHA, PLLMOD_OPT_MAX_ALPHA,
        PLLMOD_OPT_LBFGSB_BOUND_BOTH);
  if (which & PLLMOD_OPT_PARAM_FREE_RATES)
    for (unsigned int i = 0; i < K; ++i)
      add(part->rates[i], PLLMOD_OPT_MIN_RATE, PLLMOD_OPT_MAX_RATE, PLLMOD_OPT_LBFGSB_BOUND_BOTH);
  if (which & PLLMOD_OPT_PARAM_RATE_WEIGHTS)
  {
    const double * w = part->rate_weights;
    o->highest_weight_state = K - 1;
    for (unsigned int i = 1; i < K; ++i)
      if (w[i] > w[o->highest_weight_state]) o->highest_weight_state = i;
    for (unsigned int i = 0; i < K; ++i)
      if (i != o->highest_weight_state)
        add(w[i] / w[o->highest_weight_state], PLLMOD_OPT_MIN_RATE_WEIGHT,
            PLLMOD_OPT_MAX_RATE_WEIGHT, PLLMOD_OPT_LBFGSB_BOUND_BOTH);
  }
  if (which & PLLMOD_OPT_PARAM_BRANCHES_SINGLE)
    add(o->lk_params.branch_lengths[0], PLLMOD_OPT_MIN_BRANCH_LEN, PLLMOD_OPT_MAX_BRANCH_LEN,
        PLLMOD_OPT_LBFGSB_BOUND_LOWER);
  if (which & PLLMOD_OPT_PARAM_BRANCHES_ALL)
  {
    const unsigned int nb = branch_count(o);
    for (unsigned int i = 0; i < nb; ++i)
      add(o->lk_params.branch_lengths[i], PLLMOD_OPT_MIN_BRANCH_LEN, PLLMOD_OPT_MAX_BRANCH_LEN,
          PLLMOD_OPT_LBFGSB_BOUND_LOWER);
  }

  /* NOTE the reference packs "all branches" before "single branch" in set_x_to_parameters but the
   * two are mutually exclusive (assert above), so the order here is immaterial. */
  double score = pllmod_opt_minimize_lbfgsb(x.data(), lo.data(), hi.data(), kind.data(),
                                            (unsigned int)x.size(), o->factr, o->pgtol, o,
                                            negative_loglh);
  if (score != score)
  {
    score = -INFINITY;
    if (!pll_errno) pllmod_set_error(PLLMOD_OPT_ERROR_LBFGSB_UNKNOWN, "Unknown LBFGSB error");
  }
  return score;
}
