/*
 * algo_search.cpp — subtree pruning and regrafting over device-resident CLVs (SURVEY.md §8f rank 1,
 * BASELINE.json configs[4]).
 *
 *   pllmod_utree_prune / _regraft / _nodes_at_node_dist     src/tree/utree_operations.c:184-250,389-503
 *   pllmod_utree_spr / pllmod_tree_rollback (SPR)          src/tree/pll_tree.c:159-190,253-288,1894-1915
 *                                                          (libpll utree_moves.c:119-260 underneath)
 *   pllmod_algo_spr_round                                  src/algorithm/algo_search.c:603-1484
 *
 * Every candidate regraft position costs: 2 P-matrices, a partial traversal of 1-3 CLV operations
 * and one edge logL (fast mode), plus a 3-branch Newton-Raphson smoothing (thorough mode).  All of it
 * runs through the same single-launch operation-list kernel as a full traversal, so a candidate is
 * ~4 kernel launches and one host wait; no CLV ever leaves HBM and the tree surgery itself is pure
 * pointer work on the host's pll_unode_t graph.
 *
 * The control flow (candidate order, acceptance thresholds 1e-6 / 0.01, cutoff statistics, rollback
 * bookkeeping) follows the reference step by step so that a round visits the same topologies and
 * ends on the same tree; topological constraints are honoured through pllmod_treeinfo_constraint_check_spr
 * (the legacy group vector; split-based constraint trees are not built, treeinfo.cpp).
 */
#include "internal.h"
#include "trav_schedule.h"

#include <cassert>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

/* ---- tree surgery ----------------------------------------------------------------------------- */

extern "C" pll_unode_t * pllmod_utree_prune(pll_unode_t * edge)
{
  if (!edge->next)
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_SPR_INVALID_NODE, "Attempting to prune a tip node");
    return nullptr;
  }
  /* join the two neighbours; the joined branch keeps the p-matrix slot of the first one */
  pll_unode_t * a = edge->next->back;
  pll_unode_t * b = edge->next->next->back;
  pllmod_utree_connect_nodes(a, b, a->length + b->length);
  edge->next->back = edge->next->next->back = nullptr;
  return a;
}

extern "C" int pllmod_utree_regraft(pll_unode_t * edge, pll_unode_t * tree)
{
  if (!edge->next)
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_SPR_INVALID_NODE, "Attempting to regraft a tip node");
    return PLL_FAILURE;
  }
  if (edge->next->back || edge->next->next->back)
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_SPR_INVALID_NODE, "Attempting to regraft a connected node");
    return PLL_FAILURE;
  }
  /* split the target branch in two halves around the dangling inner node */
  pll_unode_t * far_end = tree->back;
  const double half = tree->length / 2;
  pllmod_utree_connect_nodes(tree, edge->next, half);
  pllmod_utree_connect_nodes(edge->next->next, far_end, half);
  return PLL_SUCCESS;
}

static void link_nodes(pll_unode_t * a, pll_unode_t * b, double length, unsigned int pmatrix_index)
{
  a->back = b;
  b->back = a;
  a->length = b->length = length;
  a->pmatrix_index = b->pmatrix_index = pmatrix_index;
}

extern "C" int pllmod_utree_spr(pll_unode_t * p, pll_unode_t * r, pll_tree_rollback_t * rb)
{
  if (pllmod_utree_is_tip(p))
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_SPR_INVALID_NODE, "Attempting to prune a leaf branch");
    return PLL_FAILURE;
  }
  if (rb)
  {
    rb->rearrange_type = PLLMOD_TREE_REARRANGE_SPR;
    rb->rooted = 0;
    rb->SPR.prune_edge = p;
    rb->SPR.regraft_edge = p->next->back;
    rb->SPR.prune_bl = p->length;
    rb->SPR.prune_left_bl = p->next->length;
    rb->SPR.prune_right_bl = p->next->next->length;
    rb->SPR.regraft_bl = r->length;
  }
  /* libpll's pll_utree_spr (utree_moves.c:119-260): note the p-matrix slot assignment differs from
   * prune + regraft above (the far half keeps the slot of p->next->next) */
  if (r == p || r == p->back || r == p->next || r == p->next->back || r == p->next->next ||
      r == p->next->next->back)
  {
    pllmod_set_error(PLL_ERROR_SPR_NOCHANGE, "Proposed move yields the same tree");
    return PLL_FAILURE;
  }
  pll_unode_t * u = p->next->back;
  pll_unode_t * v = p->next->next->back;
  link_nodes(u, v, u->length + v->length, u->pmatrix_index);
  p->next->back = p->next->next->back = nullptr;
  const double half = r->length / 2;
  link_nodes(r->back, p->next->next, half, p->next->next->pmatrix_index);
  link_nodes(r, p->next, half, r->pmatrix_index);
  return PLL_SUCCESS;
}

extern "C" int pllmod_tree_rollback(pll_tree_rollback_t * rb)
{
  if (rb->rearrange_type != PLLMOD_TREE_REARRANGE_SPR || rb->rooted)
  {
    pllmod_set_error(PLLMOD_ERROR_NOT_IMPLEMENTED, "Only unrooted SPR moves can be rolled back");
    return PLL_FAILURE;
  }
  pll_unode_t * p = (pll_unode_t *)rb->SPR.prune_edge;
  pll_unode_t * r = (pll_unode_t *)rb->SPR.regraft_edge;
  pll_unode_t * z1 = p->next->back;
  pll_unode_t * z2 = r->back;
  if (!pllmod_utree_spr(p, r, nullptr)) return PLL_FAILURE;
  pllmod_utree_set_length(z1, rb->SPR.regraft_bl);
  pllmod_utree_set_length(p, rb->SPR.prune_bl);
  pllmod_utree_set_length(r, rb->SPR.prune_left_bl);
  pllmod_utree_set_length(z2, rb->SPR.prune_right_bl);
  return PLL_SUCCESS;
}

static void collect_at_distance(pll_unode_t * node, pll_unode_t ** out, unsigned int * n,
                                unsigned int dmin, unsigned int dmax, unsigned int depth)
{
  if (depth >= dmin && depth <= dmax) out[(*n)++] = node;
  if (depth >= dmax || !node->next) return;
  collect_at_distance(node->next->back, out, n, dmin, dmax, depth + 1);
  collect_at_distance(node->next->next->back, out, n, dmin, dmax, depth + 1);
}

extern "C" int pllmod_utree_nodes_at_node_dist(pll_unode_t * node, pll_unode_t ** outbuffer,
                                               unsigned int * node_count, unsigned int min_distance,
                                               unsigned int max_distance)
{
  if (!node->next)
  {
    pllmod_set_error(PLLMOD_ERROR_INVALID_NODE_TYPE,
                     "Internal node expected, but tip node was provided");
    return PLL_FAILURE;
  }
  if (max_distance < min_distance)
  {
    pllmod_set_error(PLLMOD_ERROR_INVALID_RANGE,
                     "Invalid distance range: %d..%d (max_distance < min_distance)", min_distance,
                     max_distance);
    return PLL_FAILURE;
  }
  *node_count = 0;
  collect_at_distance(node, outbuffer, node_count, min_distance, max_distance, 0);
  return PLL_SUCCESS;
}

/* ---- search state ----------------------------------------------------------------------------- */
namespace
{
struct SearchParams
{
  bool thorough;
  unsigned int radius_min, radius_max, ntopol_keep;
  double bl_min, bl_max;
  int smoothings, brlen_opt_method;
  double lh_epsilon_triplet;
  bool fast_clv_updates;
  unsigned int sets; /* branch-length sets: partitions if unlinked, else 1 */
};

/* one scored (prune node, best regraft edge) pair with the three branch-length sets around it */
struct Candidate
{
  pll_unode_t * p_node = nullptr;
  pll_unode_t * r_node = nullptr;
  std::vector<double> b1, b2, b3;
  double lh = 0.0;
  unsigned int rollback_num = 0;
};

/* ring buffer of applied moves (algo_search.c:108-190) */
struct RollbackRing
{
  std::vector<pll_tree_rollback_t> slot;
  size_t current = 0;
  unsigned int round = 0;
  explicit RollbackRing(size_t n) : slot(n) { if (n) memset(slot.data(), 0, n * sizeof(pll_tree_rollback_t)); }
  pll_tree_rollback_t * at_current() { return slot.data() + current; } /* size 0: never dereferenced */
  pll_tree_rollback_t * next()
  {
    if (current + 1 < slot.size())
      ++current;
    else
    {
      ++round;
      current = 0;
    }
    return slot.data() + current;
  }
  pll_tree_rollback_t * prev()
  {
    if (current > 0)
      --current;
    else if (round > 0 && !slot.empty())
    {
      --round;
      current = slot.size() - 1;
    }
    else
      return nullptr;
    return slot.data() + current;
  }
  int abspos() const { return (int)(slot.size() * round + current); }
};

/* candidates that did not improve the tree, best first (algo_search.c:195-330) */
struct TopList
{
  std::vector<Candidate> list;
  explicit TopList(size_t n) : list(n) {}
  void save(const Candidate & c)
  {
    const size_t n = list.size();
    size_t idx = 0;
    while (idx < n && list[idx].p_node && c.lh < list[idx].lh) ++idx;
    if (idx >= n) return;
    for (size_t j = n - 1; j > idx; --j) list[j] = list[j - 1];
    list[idx] = c;
  }
  int next_index(unsigned int rollback_num, int cur) const
  {
    do
    {
      ++cur;
      if (cur >= (int)list.size() || !list[cur].p_node) return -1;
    } while (list[cur].rollback_num != rollback_num);
    return cur;
  }
};

double optimize_bl(pll_unode_t * node, pllmod_treeinfo_t * ti, const SearchParams & prm, int radius,
                   double lh_epsilon, double smooth_factor)
{
  const int smoothings = (int)round(smooth_factor * prm.smoothings);
  const double nll = pllmod_opt_optimize_branch_lengths_local_multi(
      ti->partitions, ti->partition_count, node, ti->param_indices, ti->deriv_precomp,
      ti->branch_lengths, ti->brlen_scalers, prm.bl_min, prm.bl_max, lh_epsilon, smoothings, radius,
      1, prm.brlen_opt_method, ti->brlen_linkage, ti->parallel_context, ti->parallel_reduce_cb);
  return nll ? -nll : 0.0;
}

/* PLLB200_TRACE_SPR=1: phase times of one round on stderr */
struct PhaseClock
{
  double search_s = 0, blo_s = 0;
  unsigned long long search_launches = 0, blo_launches = 0, blo_calls = 0;
  static double now()
  {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }
};
thread_local PhaseClock g_phase;

double optimize_bl_all(pllmod_treeinfo_t * ti, const SearchParams & prm, double lh_epsilon,
                       double smooth_factor)
{
  const double t0 = PhaseClock::now();
  const unsigned long long l0 = g_kernel_launches;
  pllmod_treeinfo_compute_loglh(ti, 0);
  const double r = optimize_bl(ti->root, ti, prm, PLLMOD_OPT_BRLEN_OPTIMIZE_ALL, lh_epsilon, smooth_factor);
  g_phase.blo_s += PhaseClock::now() - t0;
  g_phase.blo_launches += g_kernel_launches - l0;
  ++g_phase.blo_calls;
  return r;
}

/* clamp the length(s) of one branch into [bl_min, bl_max] (algo_search.c:382-424) */
void clamp_branch(pllmod_treeinfo_t * ti, pll_unode_t * node, double bl_min, double bl_max)
{
  const unsigned int m = node->pmatrix_index;
  if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
  {
    for (unsigned int p = 0; p < ti->partition_count; ++p)
    {
      if (!ti->partitions[p]) continue;
      const double l = ti->branch_lengths[p][m];
      if (l < bl_min || l > bl_max)
      {
        pllmod_treeinfo_set_branch_length_partition(ti, node, (int)p, l < bl_min ? bl_min : bl_max);
        ti->pmatrix_valid[p][m] = 0;
      }
    }
  }
  else if (node->length < bl_min || node->length > bl_max)
  {
    pllmod_treeinfo_set_branch_length(ti, node, node->length < bl_min ? bl_min : bl_max);
    pllmod_treeinfo_invalidate_pmatrix(ti, node);
  }
}

/* rebuild the (invalid) P-matrix of one edge in every local partition (algo_search.c:426-460) */
int refresh_edge_pmatrix(pllmod_treeinfo_t * ti, pll_unode_t * edge)
{
  unsigned int m = edge->pmatrix_index;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!ti->partitions[p] || ti->pmatrix_valid[p][m]) continue;
    double l = ti->branch_lengths[p][m];
    if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_SCALED) l *= ti->brlen_scalers[p];
    if (!pll_update_prob_matrices(ti->partitions[p], ti->param_indices[p], &m, &l, 1))
      return PLL_FAILURE;
    ti->pmatrix_valid[p][m] = 1;
  }
  return PLL_SUCCESS;
}

pll_unode_t * prune_with_lengths(pllmod_treeinfo_t * ti, const SearchParams & prm, pll_unode_t * edge)
{
  std::vector<double> a(prm.sets), b(prm.sets);
  if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
  {
    pllmod_treeinfo_get_branch_length_all(ti, edge->next->back, a.data());
    pllmod_treeinfo_get_branch_length_all(ti, edge->next->next->back, b.data());
    for (unsigned int i = 0; i < prm.sets; ++i) a[i] += b[i];
  }
  else
    a[0] = edge->next->next->back->length + edge->next->back->length;
  pll_unode_t * joined = pllmod_utree_prune(edge);
  if (joined) pllmod_treeinfo_set_branch_length_all(ti, joined, a.data());
  return joined;
}

int regraft_with_lengths(pllmod_treeinfo_t * ti, const SearchParams & prm, pll_unode_t * p_edge,
                         pll_unode_t * r_edge)
{
  std::vector<double> h(prm.sets);
  if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
    pllmod_treeinfo_get_branch_length_all(ti, r_edge, h.data());
  else
    h[0] = r_edge->length;
  for (double & x : h) x /= 2.;
  const int rc = pllmod_utree_regraft(p_edge, r_edge);
  pllmod_treeinfo_set_branch_length_all(ti, p_edge->next->back, h.data());
  pllmod_treeinfo_set_branch_length_all(ti, p_edge->next->next->back, h.data());
  return rc;
}

int apply_spr(pllmod_treeinfo_t * ti, const SearchParams & prm, pll_unode_t * p_edge,
              pll_unode_t * r_edge, pll_tree_rollback_t * rb)
{
  if (pllmod_utree_is_tip(p_edge))
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_SPR_INVALID_NODE, "Attempting to prune a leaf branch");
    return PLL_FAILURE;
  }
  if (rb)
  {
    rb->rearrange_type = PLLMOD_TREE_REARRANGE_SPR;
    rb->rooted = 0;
    rb->SPR.prune_edge = p_edge;
    rb->SPR.regraft_edge = p_edge->next->back;
    rb->SPR.prune_bl = p_edge->length;
    rb->SPR.prune_left_bl = p_edge->next->length;
    rb->SPR.prune_right_bl = p_edge->next->next->length;
    rb->SPR.regraft_bl = r_edge->length;
  }
  prune_with_lengths(ti, prm, p_edge);
  return regraft_with_lengths(ti, prm, p_edge, r_edge);
}

void get3(pllmod_treeinfo_t * ti, pll_unode_t * e, std::vector<double> & x1, std::vector<double> & x2,
          std::vector<double> & x3)
{
  pllmod_treeinfo_get_branch_length_all(ti, e, x1.data());
  pllmod_treeinfo_get_branch_length_all(ti, e->next, x2.data());
  pllmod_treeinfo_get_branch_length_all(ti, e->next->next, x3.data());
}
void set3(pllmod_treeinfo_t * ti, pll_unode_t * e, const std::vector<double> & x1,
          const std::vector<double> & x2, const std::vector<double> & x3)
{
  pllmod_treeinfo_set_branch_length_all(ti, e, x1.data());
  pllmod_treeinfo_set_branch_length_all(ti, e->next, x2.data());
  pllmod_treeinfo_set_branch_length_all(ti, e->next->next, x3.data());
}
void invalidate3(pllmod_treeinfo_t * ti, pll_unode_t * e)
{
  pllmod_treeinfo_invalidate_pmatrix(ti, e);
  pllmod_treeinfo_invalidate_pmatrix(ti, e->next);
  pllmod_treeinfo_invalidate_pmatrix(ti, e->next->next);
}

/* ---- batched regraft scoring (fast mode) ------------------------------------------------------------
 * The reference scores the candidates of one pruned subtree one after the other: regraft, two P-matrices, a
 * partial traversal that re-orients one to three CLVs, one edge logL, prune again (algo_search.c:727-860) —
 * here ~5 launches and a host wait each.  But all candidates share everything except the branch they split:
 * with the tree rooted at the pruning point every CLV already points TOWARDS it, and the CLV pointing AWAY from
 * it at a branch of depth d follows from the one of its parent branch and one sibling CLV.  One depth-first walk
 * to radius_max therefore emits, per branch, one ordinary operation (the away-CLV, into a scratch row per depth)
 * and one SCORE operation (k_traversal_dna4.cu: new inner node between the two halves of the branch + edge logL
 * against the pruned subtree, nothing stored), and the whole list is ONE launch of the site-column kernel plus
 * one launch for all half-length P-matrices.  Scores are computed for every branch within radius_max; the
 * reference's breadth-first loop with its cutoff rule is then replayed on the host over these numbers, so the
 * same candidates are visited in the same order and the same statistics accumulate. */
struct BatchScores
{
  std::vector<pll_unode_t *> edge; /* candidate branch (the subnode the reference's queue holds) */
  std::vector<double> score;
  double lookup(const pll_unode_t * e, bool & found) const
  {
    for (size_t i = 0; i < edge.size(); ++i)
      if (edge[i] == e)
      {
        found = true;
        return score[i];
      }
    found = false;
    return 0.0;
  }
};

struct BatchBuilder
{
  pllmod_treeinfo_t * ti;
  const SearchParams * prm;
  unsigned int nodes, scale_buffers, prob_matrices;
  std::vector<pll_operation_t> ops;
  std::vector<unsigned char> kind;
  std::vector<unsigned int> cand;
  std::vector<unsigned int> mat_idx;
  std::vector<double> mat_len;
  BatchScores * out;

  struct Src
  {
    unsigned int clv;
    int scaler;
    unsigned int matrix;
  };
  static Src real(const pll_unode_t * n, unsigned int matrix) { return Src{n->clv_index, n->scaler_index, matrix}; }

  double branch_length(const pll_unode_t * e) const
  {
    return ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED ? ti->branch_lengths[0][e->pmatrix_index] : e->length;
  }

  /* far: the subnode at the far end of a branch of depth `depth`; up: what enters its near node from the
   * pruning point's side, sib: the near node's other subtree */
  void walk(pll_unode_t * far, unsigned int depth, const Src & up, const Src & sib)
  {
    pll_operation_t o;
    o.parent_clv_index = nodes + depth;
    o.parent_scaler_index = (int)(scale_buffers + depth);
    o.child1_clv_index = up.clv;
    o.child1_scaler_index = up.scaler;
    o.child1_matrix_index = up.matrix;
    o.child2_clv_index = sib.clv;
    o.child2_scaler_index = sib.scaler;
    o.child2_matrix_index = sib.matrix;
    ops.push_back(o);
    kind.push_back(0);
    cand.push_back(0);
    if (depth >= prm->radius_min)
    {
      /* both halves get half the branch, clamped (regraft_with_lengths + clamp_branch) */
      double half = branch_length(far) / 2.;
      if (half < prm->bl_min) half = prm->bl_min;
      if (half > prm->bl_max) half = prm->bl_max;
      if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_SCALED) half *= ti->brlen_scalers[0];
      const unsigned int m = prob_matrices + (unsigned int)mat_idx.size();
      mat_idx.push_back(m);
      mat_len.push_back(half);
      pll_operation_t sop;
      sop.parent_clv_index = nodes; /* scratch row 0: written by nobody, read by nobody */
      sop.parent_scaler_index = -1;
      sop.child1_clv_index = far->clv_index; /* points towards the pruning point: valid */
      sop.child1_scaler_index = far->scaler_index;
      sop.child1_matrix_index = m;
      sop.child2_clv_index = nodes + depth; /* the away-CLV just computed: forwarded in registers */
      sop.child2_scaler_index = (int)(scale_buffers + depth);
      sop.child2_matrix_index = m;
      ops.push_back(sop);
      kind.push_back(LSPECIAL_SCORE);
      cand.push_back((unsigned int)out->edge.size());
      out->edge.push_back(far);
    }
    if (depth >= prm->radius_max || !far->next) return;
    pll_unode_t * c1 = far->next;
    pll_unode_t * c2 = far->next->next;
    const Src here{nodes + depth, (int)(scale_buffers + depth), far->pmatrix_index};
    walk(c1->back, depth + 1, here, real(c2->back, c2->pmatrix_index));
    walk(c2->back, depth + 1, here, real(c1->back, c1->pmatrix_index));
  }
};

/* true: `scores` holds the logL of every candidate within radius_max of the (already pruned) subtree behind
 * p_edge; false: this case is not eligible (the caller scores sequentially) or an error occurred (pll_errno) */
static bool score_candidates_batched(pllmod_treeinfo_t * ti, const SearchParams & prm, pll_unode_t * p_edge,
                                     BatchScores & scores)
{
  if (!g_opt_spr_batch || prm.thorough || ti->partition_count != 1 || !ti->partitions[0]) return false;
  pll_partition_t * part = ti->partitions[0];
  PartCtx * c = ctx_of(part);
  if (c->magic != PLLB200_MAGIC || !lean_supported(c) || (part->attributes & PLL_ATTRIB_AB_MASK)) return false;
  if (c->scratch_rows < prm.radius_max + 2 || !part->scale_buffers || prm.radius_min < 1) return false;
  for (unsigned int k = 0; k < part->rate_cats; ++k)
    if (part->prop_invar[ti->param_indices[0][k]] > 0.0) return false;
  pll_unode_t * R = ti->root; /* the joined branch left behind by the pruning */
  /* every CLV must point towards the pruning point (the two at the joined branch were invalidated) */
  if (!pllmod_treeinfo_update_partials(ti, 1)) return false;
  BatchBuilder b;
  b.ti = ti;
  b.prm = &prm;
  b.nodes = part->nodes;
  b.scale_buffers = part->scale_buffers;
  b.prob_matrices = part->prob_matrices;
  b.out = &scores;
  scores.edge.clear();
  /* SETUP: the pruned subtree's CLV over the prune branch (p_edge keeps its length in every candidate) */
  {
    pll_operation_t o;
    o.parent_clv_index = part->nodes;
    o.parent_scaler_index = -1;
    o.child1_clv_index = p_edge->back->clv_index; /* SETUP uses child 2 only: child 1 names no scaler, so that */
    o.child1_scaler_index = -1;                   /* the subtree's counts are not taken twice */
    o.child1_matrix_index = p_edge->pmatrix_index;
    o.child2_clv_index = p_edge->back->clv_index;
    o.child2_scaler_index = p_edge->back->scaler_index;
    o.child2_matrix_index = p_edge->pmatrix_index;
    b.ops.push_back(o);
    b.kind.push_back(LSPECIAL_SETUP);
    b.cand.push_back(0);
  }
  for (int side = 0; side < 2; ++side)
  {
    pll_unode_t * top = side ? R->back : R;
    if (!top->next) continue;
    pll_unode_t * n1 = top->next;
    pll_unode_t * n2 = top->next->next;
    const BatchBuilder::Src up = BatchBuilder::real(top->back, top->pmatrix_index);
    b.walk(n1->back, 1, up, BatchBuilder::real(n2->back, n2->pmatrix_index));
    b.walk(n2->back, 1, up, BatchBuilder::real(n1->back, n1->pmatrix_index));
  }
  const unsigned int ncand = (unsigned int)scores.edge.size();
  if (!ncand || ncand > c->scratch_mats || ncand > PLLB200_SCORE_MAX || b.ops.size() > 512) return false;
  if (!pll_update_prob_matrices(part, ti->param_indices[0], b.mat_idx.data(), b.mat_len.data(), ncand)) return false;
  scores.score.assign(ncand, 0.0);
  if (!launch_trav_dna4_score(c, b.ops.data(), b.kind.data(), b.cand.data(), (unsigned int)b.ops.size(), ncand,
                              ti->param_indices[0], scores.score.data()))
    return false;
  if (ti->parallel_reduce_cb)
    ti->parallel_reduce_cb(ti->parallel_context, scores.score.data(), ncand, PLLMOD_COMMON_REDUCE_SUM);
  return true;
}

/* Try every regraft edge within the radius window for the subtree behind `cand.p_node`, remember the
 * best one, and put the subtree back where it was (algo_search.c:603-899). */
int best_reinsert_edge(pllmod_treeinfo_t * ti, Candidate & cand, cutoff_info_t * cutoff,
                       const SearchParams & prm)
{
  pll_unode_t * p_edge = cand.p_node;
  const size_t total_edges = ti->tree->edge_count;
  const unsigned int S = prm.sets;
  std::vector<double> z1(S), z2(S), z3(S), b1(S), b2(S), b3(S), regraft_len(S);
  cand.r_node = nullptr;
  cand.lh = PLLMOD_OPT_LNL_UNLIKELY;

  get3(ti, p_edge, z1, z2, z3);
  pllmod_treeinfo_set_root(ti, p_edge);
  double loglh;
  if (prm.fast_clv_updates)
  {
    pllmod_treeinfo_invalidate_clv(ti, p_edge);
    loglh = pllmod_treeinfo_compute_loglh_flex(ti, 1, 0);
  }
  else
    loglh = pllmod_treeinfo_compute_loglh(ti, 0);

  /* topological constraint (algo_search.c:661-662,736-742): candidates that contradict it are skipped */
  pllmod_treeinfo_constraint_update_splits(ti);
  const int check_cons = pllmod_treeinfo_constraint_subtree_affected(ti, p_edge);

  /* prune; the joined branch becomes the temporary root edge */
  pll_unode_t * orig_prune_edge = prune_with_lengths(ti, prm, p_edge);
  if (!orig_prune_edge) return PLL_FAILURE;
  clamp_branch(ti, orig_prune_edge, prm.bl_min, prm.bl_max);
  pllmod_treeinfo_set_root(ti, orig_prune_edge);
  pllmod_treeinfo_invalidate_clv(ti, orig_prune_edge);
  pllmod_treeinfo_invalidate_clv(ti, orig_prune_edge->back);
  pllmod_treeinfo_invalidate_pmatrix(ti, orig_prune_edge);
  if (!refresh_edge_pmatrix(ti, orig_prune_edge)) return PLL_FAILURE;

  /* fast mode, one Gamma-4 DNA partition: every candidate within radius_max is scored by one launch */
  BatchScores scores;
  pll_errno = 0;
  bool batched = score_candidates_batched(ti, prm, p_edge, scores);
  if (!batched && pll_errno) return PLL_FAILURE;
  /* debugging aid (PLLB200_SPR_VERIFY=1): score sequentially as well and report every disagreement */
  static const bool verify = getenv("PLLB200_SPR_VERIFY") != nullptr;
  const bool have_scores = batched;
  if (verify) batched = false;

  /* breadth-first list of regraft candidates, seeded at distance radius_min on both sides */
  std::vector<pll_unode_t *> queue(total_edges + 2, nullptr);
  std::vector<unsigned int> dist(total_edges + 2, 0);
  unsigned int queued = 0, n = 0;
  int ok = pllmod_utree_nodes_at_node_dist(ti->root, &queue[queued], &n, prm.radius_min, prm.radius_min);
  queued += n;
  if (!pllmod_utree_is_tip(ti->root->back))
  {
    ok &= pllmod_utree_nodes_at_node_dist(ti->root->back, &queue[queued], &n, prm.radius_min,
                                          prm.radius_min);
    queued += n;
  }
  if (!ok) return PLL_FAILURE;
  for (unsigned int i = 0; i < queued; ++i) dist[i] = prm.radius_min;

  for (unsigned int j = 0; queue[j] != nullptr; ++j)
  {
    pll_unode_t * r_edge = queue[j];
    if (r_edge == orig_prune_edge || r_edge == orig_prune_edge->back) continue;
    if (check_cons && !pllmod_treeinfo_constraint_check_spr(ti, p_edge, r_edge)) continue;
    const unsigned int r_dist = dist[j];
    if (batched)
    {
      bool found = false;
      loglh = scores.lookup(r_edge, found);
      if (!found)
      {
        pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE, "Regraft candidate outside the scored neighbourhood");
        return PLL_FAILURE;
      }
      if (loglh > cand.lh)
      {
        cand.lh = loglh;
        cand.r_node = r_edge;
      }
    }
    else
    {
      pllmod_treeinfo_get_branch_length_all(ti, r_edge, regraft_len.data());

      if (!regraft_with_lengths(ti, prm, p_edge, r_edge)) return PLL_FAILURE;
      pllmod_treeinfo_set_root(ti, p_edge);
      pllmod_treeinfo_invalidate_clv(ti, p_edge);
      get3(ti, p_edge, b1, b2, b3);
      clamp_branch(ti, p_edge->next, prm.bl_min, prm.bl_max);
      clamp_branch(ti, p_edge->next->next, prm.bl_min, prm.bl_max);
      pllmod_treeinfo_invalidate_pmatrix(ti, p_edge->next);
      pllmod_treeinfo_invalidate_pmatrix(ti, p_edge->next->next);
      if (!refresh_edge_pmatrix(ti, p_edge->next) || !refresh_edge_pmatrix(ti, p_edge->next->next))
        return PLL_FAILURE;

      /* score: partial traversal (+ triplet optimisation in thorough mode) */
      loglh = pllmod_treeinfo_compute_loglh_flex(ti, 1, 0);
      if (prm.thorough)
      {
        loglh = optimize_bl(p_edge, ti, prm, 1, prm.lh_epsilon_triplet, 1.0);
        if (!loglh) return PLL_FAILURE;
      }
      if (verify && have_scores)
      {
        bool found = false;
        const double bs = scores.lookup(r_edge, found);
        if (!found || fabs(bs - loglh) > 1e-9 * fabs(loglh))
          fprintf(stderr, "[spr verify] prune clv %u candidate clv %u (back %u) dist %u: sequential %.12g batched %.12g%s\n",
                  p_edge->clv_index, r_edge->clv_index, r_edge->back->clv_index, r_dist, loglh, bs, found ? "" : " (not scored)");
      }
      if (loglh > cand.lh)
      {
        cand.lh = loglh;
        cand.r_node = r_edge;
        get3(ti, p_edge, cand.b1, cand.b2, cand.b3);
      }

      /* undo: original lengths, prune again, restore the split branch */
      set3(ti, p_edge, b1, b2, b3);
      invalidate3(ti, p_edge);
      pll_unode_t * rejoined = pllmod_utree_prune(p_edge);
      pllmod_treeinfo_set_branch_length_all(ti, rejoined, regraft_len.data());
      pllmod_treeinfo_invalidate_pmatrix(ti, rejoined);
      if (!refresh_edge_pmatrix(ti, p_edge) || !refresh_edge_pmatrix(ti, rejoined)) return PLL_FAILURE;
    }

    bool descend = r_dist < prm.radius_max;
    if (cutoff && loglh < cutoff->lh_start)
    {
      cutoff->lh_dec_count++;
      cutoff->lh_dec_sum += cutoff->lh_start - loglh;
      descend = descend && (cutoff->lh_start - loglh) < cutoff->lh_cutoff;
    }
    if (r_edge->next && descend)
    {
      queue[queued] = r_edge->next->back;
      queue[queued + 1] = r_edge->next->next->back;
      dist[queued] = dist[queued + 1] = r_dist + 1;
      queued += 2;
    }
    assert(j + 1 < total_edges);
  }

  if (!prm.fast_clv_updates) pllmod_treeinfo_set_root(ti, orig_prune_edge);
  /* back into the original position with the original lengths */
  if (!pllmod_utree_regraft(p_edge, orig_prune_edge) && !(pll_errno & PLLMOD_TREE_ERROR_SPR_MASK))
    return PLL_FAILURE;
  set3(ti, p_edge, z1, z2, z3);
  invalidate3(ti, p_edge);
  if (prm.fast_clv_updates)
  {
    if (!refresh_edge_pmatrix(ti, p_edge) || !refresh_edge_pmatrix(ti, p_edge->next) ||
        !refresh_edge_pmatrix(ti, p_edge->next->next))
      return PLL_FAILURE;
    pllmod_treeinfo_set_root(ti, p_edge);
    pllmod_treeinfo_invalidate_clv(ti, p_edge);
    loglh = pllmod_treeinfo_compute_loglh_flex(ti, 1, 0);
  }
  (void)loglh;
  return PLL_SUCCESS;
}

/* one pass over `nodes`: apply improving moves, file the others in the top list
 * (algo_search.c:901-1037).  Returns the last logL, 0 on error. */
double reinsert_nodes(pllmod_treeinfo_t * ti, pll_unode_t ** nodes, int node_count,
                      RollbackRing & ring, TopList & top, cutoff_info_t * cutoff,
                      const SearchParams & prm)
{
  double loglh = pllmod_treeinfo_compute_loglh(ti, 0);
  double best_lh = loglh;
  Candidate cand;
  cand.b1.assign(prm.sets, 0.0);
  cand.b2.assign(prm.sets, 0.0);
  cand.b3.assign(prm.sets, 0.0);
  pll_tree_rollback_t * rb = ring.at_current();

  for (int i = 0; i < node_count; ++i)
  {
    pll_unode_t * p_edge = nodes[i];
    /* the rest of the tree would be a cherry: nothing to regraft into */
    if (pllmod_utree_is_tip(p_edge->next->back) && pllmod_utree_is_tip(p_edge->next->next->back))
      continue;
    cand.p_node = p_edge;
    if (cutoff) cutoff->lh_start = best_lh;
    if (!best_reinsert_edge(ti, cand, cutoff, prm)) return 0;

    pll_unode_t * best_r = cand.r_node;
    if (!best_r || best_r == p_edge || best_r == p_edge->back || best_r->back == p_edge) continue;

    if (cand.lh - best_lh > 1e-6)
    {
      pll_unode_t * orig_prune_edge = p_edge->next->back;
      if (!apply_spr(ti, prm, p_edge, best_r, rb)) return 0;
      clamp_branch(ti, orig_prune_edge, prm.bl_min, prm.bl_max);
      if (prm.fast_clv_updates)
      {
        pllmod_treeinfo_invalidate_pmatrix(ti, orig_prune_edge);
        if (!refresh_edge_pmatrix(ti, orig_prune_edge)) return 0;
      }
      rb = ring.next();
      if (prm.thorough)
        set3(ti, p_edge, cand.b1, cand.b2, cand.b3);
      else
      {
        clamp_branch(ti, p_edge->next, prm.bl_min, prm.bl_max);
        clamp_branch(ti, p_edge->next->next, prm.bl_min, prm.bl_max);
      }
      if (prm.fast_clv_updates)
      {
        invalidate3(ti, p_edge);
        if (!refresh_edge_pmatrix(ti, p_edge) || !refresh_edge_pmatrix(ti, p_edge->next) ||
            !refresh_edge_pmatrix(ti, p_edge->next->next))
          return 0;
        pllmod_treeinfo_invalidate_clv(ti, p_edge);
        loglh = pllmod_treeinfo_compute_loglh_flex(ti, 1, 0);
      }
      best_lh = cand.lh;
    }
    else
    {
      cand.rollback_num = (unsigned int)ring.abspos();
      top.save(cand);
      loglh = cand.lh;
    }
  }
  return loglh;
}

void collect_inner_subnodes(pll_unode_t * node, std::vector<pll_unode_t *> & out)
{
  if (!node->next) return;
  collect_inner_subnodes(node->next->back, out);
  collect_inner_subnodes(node->next->next->back, out);
  out.push_back(node->next->next);
  out.push_back(node->next);
  out.push_back(node);
}
}  // namespace

extern "C" double pllmod_algo_spr_round(pllmod_treeinfo_t * ti, unsigned int radius_min,
                                        unsigned int radius_max, unsigned int ntopol_keep,
                                        pll_bool_t thorough, int brlen_opt_method, double bl_min,
                                        double bl_max, int smoothings, double epsilon,
                                        cutoff_info_t * cutoff, double subtree_cutoff,
                                        double lh_epsilon_brlen_triplet, pll_bool_t fast_clv_updates)
{
  SearchParams prm;
  prm.thorough = thorough != 0;
  prm.radius_min = radius_min;
  prm.radius_max = radius_max;
  prm.ntopol_keep = ntopol_keep;
  prm.bl_min = bl_min;
  prm.bl_max = bl_max;
  prm.smoothings = smoothings;
  prm.brlen_opt_method = brlen_opt_method;
  prm.lh_epsilon_triplet = lh_epsilon_brlen_triplet;
  prm.fast_clv_updates = fast_clv_updates != 0;
  const bool unlinked = ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED;
  prm.sets = unlinked ? ti->init_partition_count : 1;
  pll_errno = 0;
  /* make sure the initial topology is compatible with the constraint (algo_search.c:1113-1119) */
  if (!pllmod_treeinfo_constraint_check_current(ti))
  {
    if (pll_errno != PLLMOD_ERROR_NOT_IMPLEMENTED)
      pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE, "Constraint check failed before SPR round!");
    return 0;
  }

  RollbackRing ring(ntopol_keep);
  TopList top(prm.thorough ? ntopol_keep : ntopol_keep * 3);
  if (cutoff)
  {
    cutoff->lh_dec_count = 0;
    cutoff->lh_dec_sum = 0.;
  }

  double loglh = pllmod_treeinfo_compute_loglh(ti, 0);
  pll_unode_t * initial_root = ti->root;
  double best_lh = loglh;

  std::vector<pll_unode_t *> allnodes;
  allnodes.reserve((size_t)(ti->tip_count - 2) * 3);
  collect_inner_subnodes(ti->root->back, allnodes);
  collect_inner_subnodes(ti->root, allnodes);

  g_phase = PhaseClock();
  {
    const double t0 = PhaseClock::now();
    const unsigned long long l0 = g_kernel_launches;
    loglh = reinsert_nodes(ti, allnodes.data(), (int)allnodes.size(), ring, top, cutoff, prm);
    g_phase.search_s = PhaseClock::now() - t0;
    g_phase.search_launches = g_kernel_launches - l0;
  }
  if (!loglh) return 0;

  /* NOTE: the reference has a second, thorough pass over the best rejected subtrees in fast mode, but
   * guards it with `bestnode_list->current > 0`, a counter that is never incremented
   * (algo_search.c:219,1194) — the pass never runs there, so it does not exist here either. */
  if (prm.fast_clv_updates) pllmod_treeinfo_set_root(ti, initial_root);

  best_lh = optimize_bl_all(ti, prm, epsilon, 0.25);
  if (!best_lh) return 0;
  pllmod_treeinfo_topology_t * best_topol = pllmod_treeinfo_get_topology(ti, nullptr);
  if (!best_topol) return 0;

  /* Walk back through the applied moves; at each stage re-apply the rejected candidates recorded at
   * that stage, optimise all branches and keep the best tree (algo_search.c:1252-1408). */
  size_t undone = 0;
  int top_index = -1;
  pll_tree_rollback_t undo_move;
  memset(&undo_move, 0, sizeof(undo_move));
  while (undone < ring.slot.size())
  {
    const int stage = ring.abspos();
    top_index = top.next_index((unsigned int)stage, top_index);
    bool undo_after = false;
    if (top_index == -1)
    {
      pll_tree_rollback_t * rb = ring.prev();
      if (!rb || !rb->SPR.prune_edge) break;
      /* algo_search.c:1297-1310: a stage whose move contradicts the constraint is undone but not scored */
      pllmod_treeinfo_constraint_update_splits(ti);
      const bool skip_topol = !pllmod_treeinfo_constraint_check_spr(ti, (pll_unode_t *)rb->SPR.prune_edge,
                                                                     (pll_unode_t *)rb->SPR.regraft_edge);
      if (!pllmod_tree_rollback(rb)) goto fail;
      ++undone;
      if (skip_topol) continue;
    }
    else
    {
      if ((unsigned int)top_index > prm.ntopol_keep) continue;
      Candidate & c = top.list[top_index];
      if (!c.p_node) break;
      pllmod_treeinfo_constraint_update_splits(ti);
      if (!pllmod_treeinfo_constraint_check_spr(ti, c.p_node, c.r_node)) continue; /* algo_search.c:1339-1344 */
      if (!pllmod_utree_spr(c.p_node, c.r_node, &undo_move)) goto fail;
      clamp_branch(ti, (pll_unode_t *)undo_move.SPR.regraft_edge, prm.bl_min, prm.bl_max);
      if (thorough)
        set3(ti, c.p_node, c.b1, c.b2, c.b3);
      else
      {
        clamp_branch(ti, c.p_node, prm.bl_min, prm.bl_max);
        clamp_branch(ti, c.p_node->next, prm.bl_min, prm.bl_max);
        clamp_branch(ti, c.p_node->next->next, prm.bl_min, prm.bl_max);
      }
      undo_after = true;
    }
    {
      const double lh = optimize_bl_all(ti, prm, epsilon, 0.25);
      if (!lh) goto fail;
      if (lh - best_lh > 0.01)
      {
        best_topol = pllmod_treeinfo_get_topology(ti, best_topol);
        if (!best_topol) return 0;
        best_lh = lh;
      }
    }
    if (undo_after && !pllmod_tree_rollback(&undo_move)) goto fail;
  }

  if (cutoff) cutoff->lh_cutoff = subtree_cutoff * (cutoff->lh_dec_sum / cutoff->lh_dec_count);
  {
    const int rc = pllmod_treeinfo_set_topology(ti, best_topol);
    pllmod_treeinfo_destroy_topology(best_topol);
    best_topol = nullptr;
    if (!rc) return 0;
  }
  loglh = pllmod_treeinfo_compute_loglh(ti, 0);
  if (getenv("PLLB200_TRACE_SPR"))
    fprintf(stderr, "[spr_round] search %.3f s (%llu launches), all-branch optimisation %.3f s in %llu calls (%llu launches)\n",
            g_phase.search_s, g_phase.search_launches, g_phase.blo_s, g_phase.blo_calls, g_phase.blo_launches);
  if (fabs(loglh - best_lh) > 1e-6)
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE, "LH mismatch after SPR round: %.12f != %.12f",
                     best_lh, loglh);
    return 0;
  }
  return loglh;

fail:
  if (best_topol) pllmod_treeinfo_destroy_topology(best_topol);
  if (!pll_errno) pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE, "SPR round failed");
  return 0;
}
