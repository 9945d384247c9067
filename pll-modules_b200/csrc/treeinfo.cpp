/*
 * treeinfo.cpp — pllmod_treeinfo_*: the multi-partition tree + model container that turns a tree
 * into operation lists and tracks CLV / P-matrix validity (reference: src/tree/treeinfo.c).
 *
 * Same public struct, same validity rules, same results; what differs is how the device is fed:
 *   - invalid P-matrices of a partition are gathered and rebuilt in ONE batched call
 *     (the reference issues one pll_update_prob_matrices call per matrix, treeinfo.c:845-865);
 *   - the per-partition loop of treeinfo.c:1022-1058 only QUEUES work: CLV traversal and the edge
 *     log-likelihood of every partition go to that partition's stream, and the host waits once for
 *     all of them before the parallel_reduce_cb / sum (treeinfo.c:1061-1073);
 *   - deriv_precomp[p] is a 64-byte key for a device-resident sumtable, not a CLV-sized host buffer.
 * Topological constraints: the legacy group vector is built (set_constraint_clvmap, constraint_check_spr);
 * the split-set machinery (utree_constraint.c) is outside the hot path and not built.
 */
#include "internal.h"

#include <cassert>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

static int cb_full_traversal(pll_unode_t *) { return PLL_SUCCESS; }

/* partial traversal: descend only where some active partition has a stale CLV (treeinfo.c:38-61) */
static int cb_partial_traversal(pll_unode_t * node)
{
  if (!node->next) return PLL_FAILURE;
  pllmod_treeinfo_t * ti = (pllmod_treeinfo_t *)node->data;
  if (ti->active_partition == PLLMOD_TREEINFO_PARTITION_ALL)
  {
    for (unsigned int i = 0; i < ti->init_partition_count; ++i)
      if (!ti->clv_valid[ti->init_partition_idx[i]][node->node_index]) return PLL_SUCCESS;
    return PLL_FAILURE;
  }
  return ti->clv_valid[ti->active_partition][node->node_index] == 0;
}

static inline bool part_active(const pllmod_treeinfo_t * ti, unsigned int p)
{
  return ti->active_partition == PLLMOD_TREEINFO_PARTITION_ALL || ti->active_partition == (int)p;
}

static int check_tree(pllmod_treeinfo_t * ti, pll_utree_t * tree)
{
  if (!ti || !tree)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Parameter is NULL\n");
    return PLL_FAILURE;
  }
  if (ti->tip_count != tree->tip_count)
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE_SIZE, "Invalid tree size. Got %d instead of %d\n",
                     tree->tip_count, ti->tip_count);
    return PLL_FAILURE;
  }
  if (!tree->binary)
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE, "Binary tree expected\n");
    return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}

/* back-pointers node->data = treeinfo, subnode table, branch lengths by pmatrix index
 * (treeinfo.c:1242-1293) */
static int init_tree(pllmod_treeinfo_t * ti)
{
  pll_utree_t * tree = ti->tree;
  const unsigned int node_count = tree->tip_count + tree->inner_count;
  const unsigned int edge_count = tree->edge_count;
  ti->root = tree->vroot;
  for (unsigned int i = 0; i < node_count; ++i)
  {
    pll_unode_t * sn = tree->nodes[i];
    do
    {
      if (sn->pmatrix_index >= edge_count || sn->node_index >= ti->subnode_count)
      {
        pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE,
                         "p-matrix index (%u) or node index (%u) out of bounds: every branch needs "
                         "a unique p-matrix index in [0, branch_count)\n",
                         sn->pmatrix_index, sn->node_index);
        return PLL_FAILURE;
      }
      ti->subnodes[sn->node_index] = sn;
      ti->branch_lengths[0][sn->pmatrix_index] = sn->length;
      sn->data = ti;
      sn = sn->next;
    } while (sn && sn != tree->nodes[i]);
  }
  if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
    for (unsigned int p = 1; p < ti->partition_count; ++p)
      memcpy(ti->branch_lengths[p], ti->branch_lengths[0], sizeof(double) * edge_count);
  return PLL_SUCCESS;
}

extern "C" pllmod_treeinfo_t * pllmod_treeinfo_create(pll_unode_t * root, unsigned int tips,
                                                      unsigned int partitions, int brlen_linkage)
{
  pllmod_treeinfo_t * ti = (pllmod_treeinfo_t *)calloc(1, sizeof(pllmod_treeinfo_t));
  if (!ti)
  {
    pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for treeinfo\n");
    return nullptr;
  }
  ti->tip_count = tips;
  ti->partition_count = partitions;
  ti->brlen_linkage = brlen_linkage;
  ti->tree = pll_utree_wraptree(root, tips);
  if (!ti->tree || !check_tree(ti, ti->tree))
  {
    if (ti->tree)
    {
      free(ti->tree->nodes);
      free(ti->tree);
    }
    free(ti);
    return nullptr;
  }
  const unsigned int inner = ti->tree->inner_count;
  const unsigned int nodes = inner + tips;
  const unsigned int branches = ti->tree->edge_count;
  ti->subnode_count = tips + 3 * inner;

  ti->travbuffer = (pll_unode_t **)malloc(sizeof(pll_unode_t *) * nodes);
  ti->matrix_indices = (unsigned int *)malloc(sizeof(unsigned int) * branches);
  ti->operations = (pll_operation_t *)malloc(sizeof(pll_operation_t) * inner);
  ti->subnodes = (pll_unode_t **)calloc(ti->subnode_count, sizeof(pll_unode_t *));
  ti->partitions = (pll_partition_t **)calloc(partitions, sizeof(pll_partition_t *));
  ti->params_to_optimize = (int *)calloc(partitions, sizeof(int));
  ti->alphas = (double *)calloc(partitions, sizeof(double));
  ti->gamma_mode = (int *)calloc(partitions, sizeof(int));
  ti->param_indices = (unsigned int **)calloc(partitions, sizeof(unsigned int *));
  ti->subst_matrix_symmetries = (int **)calloc(partitions, sizeof(int *));
  ti->branch_lengths = (double **)calloc(partitions, sizeof(double *));
  ti->deriv_precomp = (double **)calloc(partitions, sizeof(double *));
  ti->clv_valid = (char **)calloc(partitions, sizeof(char *));
  ti->pmatrix_valid = (char **)calloc(partitions, sizeof(char *));
  ti->partition_loglh = (double *)calloc(partitions, sizeof(double));
  ti->init_partition_idx = (unsigned int *)calloc(partitions, sizeof(unsigned int));
  ti->init_partitions = (pll_partition_t **)calloc(partitions, sizeof(pll_partition_t *));
  ti->linked_branch_lengths = (double *)malloc(sizeof(double) * branches);
  ti->brlen_scalers =
      brlen_linkage == PLLMOD_COMMON_BRLEN_SCALED ? (double *)calloc(partitions, sizeof(double)) : nullptr;
  if (!ti->travbuffer || !ti->matrix_indices || !ti->operations || !ti->subnodes ||
      !ti->partitions || !ti->params_to_optimize || !ti->alphas || !ti->gamma_mode ||
      !ti->param_indices || !ti->subst_matrix_symmetries || !ti->branch_lengths ||
      !ti->deriv_precomp || !ti->clv_valid || !ti->pmatrix_valid || !ti->partition_loglh ||
      !ti->init_partition_idx || !ti->init_partitions || !ti->linked_branch_lengths ||
      (brlen_linkage == PLLMOD_COMMON_BRLEN_SCALED && !ti->brlen_scalers))
  {
    pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for treeinfo structures\n");
    pllmod_treeinfo_destroy(ti);
    return nullptr;
  }
  for (unsigned int p = 0; p < partitions; ++p)
  {
    ti->gamma_mode[p] = PLL_GAMMA_RATES_MEAN;
    ti->branch_lengths[p] = brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED
                                ? (double *)malloc(sizeof(double) * branches)
                                : ti->linked_branch_lengths;
    if (ti->brlen_scalers) ti->brlen_scalers[p] = 1.0;
    if (!ti->branch_lengths[p])
    {
      pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for arrays for partition %d\n", p);
      pllmod_treeinfo_destroy(ti);
      return nullptr;
    }
  }
  ti->active_partition = PLLMOD_TREEINFO_PARTITION_ALL;
  if (!init_tree(ti))
  {
    pllmod_treeinfo_destroy(ti);
    return nullptr;
  }
  return ti;
}

extern "C" int pllmod_treeinfo_set_parallel_context(pllmod_treeinfo_t * ti, void * ctx,
                                                    void (*cb)(void *, double *, size_t, int))
{
  ti->parallel_context = ctx;
  ti->parallel_reduce_cb = cb;
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_init_partition(pllmod_treeinfo_t * ti, unsigned int pidx,
                                              pll_partition_t * partition, int params_to_optimize,
                                              int gamma_mode, double alpha,
                                              const unsigned int * param_indices,
                                              const int * subst_matrix_symmetries)
{
  if (!ti)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Treeinfo structure is NULL\n");
    return PLL_FAILURE;
  }
  if (pidx >= ti->partition_count)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Partition %d is out of bounds\n", pidx);
    return PLL_FAILURE;
  }
  if (ti->partitions[pidx])
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Partition %d is already initialized\n", pidx);
    return PLL_FAILURE;
  }
  const unsigned int local = ti->init_partition_count++;
  ti->partitions[pidx] = partition;
  ti->init_partitions[local] = partition;
  ti->init_partition_idx[local] = pidx;
  ti->params_to_optimize[pidx] = params_to_optimize;
  ti->gamma_mode[pidx] = gamma_mode;
  ti->alphas[pidx] = alpha;

  const unsigned int inner = ti->tip_count - 2;
  ti->clv_valid[pidx] = (char *)calloc(inner * 3 + ti->tip_count, 1);
  ti->pmatrix_valid[pidx] = (char *)calloc(2 * ti->tip_count - 3, 1);
  ti->param_indices[pidx] = (unsigned int *)calloc(partition->rate_cats, sizeof(unsigned int));
  if (!ti->clv_valid[pidx] || !ti->pmatrix_valid[pidx] || !ti->param_indices[pidx])
  {
    pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for parameter indices\n");
    return PLL_FAILURE;
  }
  if (param_indices)
    memcpy(ti->param_indices[pidx], param_indices, sizeof(unsigned int) * partition->rate_cats);
  if (subst_matrix_symmetries)
  {
    const size_t sz = sizeof(int) * (partition->states * (partition->states - 1) / 2);
    ti->subst_matrix_symmetries[pidx] = (int *)malloc(sz);
    if (!ti->subst_matrix_symmetries[pidx])
    {
      pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for substitution scheme\n");
      return PLL_FAILURE;
    }
    memcpy(ti->subst_matrix_symmetries[pidx], subst_matrix_symmetries, sz);
  }
  /* key for the device-resident sumtable (treeinfo.c:332-349 allocates a CLV-sized host buffer;
   * released with free() like treeinfo.c:761-765) */
  ti->deriv_precomp[pidx] = (double *)pll_aligned_alloc(64, 32);
  if (!ti->deriv_precomp[pidx])
  {
    pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for derivative buffers\n");
    return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_set_active_partition(pllmod_treeinfo_t * ti, int pidx)
{
  if (pidx != PLLMOD_TREEINFO_PARTITION_ALL && pidx >= (int)ti->partition_count)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Partition %d is out of bounds\n", pidx);
    return PLL_FAILURE;
  }
  ti->active_partition = pidx;
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_set_root(pllmod_treeinfo_t * ti, pll_unode_t * root)
{
  if (!ti || !root || root->data != (void *)ti)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Invalid root node!\n");
    return PLL_FAILURE;
  }
  ti->root = pllmod_utree_is_tip(root) ? root->back : root; /* root must be inner */
  ti->tree->vroot = ti->root;
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_get_branch_length_all(const pllmod_treeinfo_t * ti,
                                                     const pll_unode_t * edge, double * lengths)
{
  const unsigned int m = edge->pmatrix_index;
  if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
  {
    for (unsigned int p = 0; p < ti->partition_count; ++p)
      if (ti->partitions[p]) *lengths++ = ti->branch_lengths[p][m];
  }
  else
    lengths[0] = ti->branch_lengths[0][m];
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_set_branch_length_all(pllmod_treeinfo_t * ti, pll_unode_t * edge,
                                                     const double * lengths)
{
  const unsigned int m = edge->pmatrix_index;
  if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED)
  {
    for (unsigned int p = 0; p < ti->partition_count; ++p)
      if (ti->partitions[p]) ti->branch_lengths[p][m] = *lengths++;
  }
  else
  {
    ti->branch_lengths[0][m] = lengths[0];
    pllmod_utree_set_length(edge, lengths[0]);
  }
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_set_branch_length(pllmod_treeinfo_t * ti, pll_unode_t * edge,
                                                 double length)
{
  assert(ti->brlen_linkage != PLLMOD_COMMON_BRLEN_UNLINKED);
  return pllmod_treeinfo_set_branch_length_all(ti, edge, &length);
}

extern "C" int pllmod_treeinfo_set_branch_length_partition(pllmod_treeinfo_t * ti,
                                                           pll_unode_t * edge, int pidx,
                                                           double length)
{
  const unsigned int m = edge->pmatrix_index;
  if (pidx != PLLMOD_TREEINFO_PARTITION_ALL && pidx >= (int)ti->partition_count)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Partition %d is out of bounds\n", pidx);
    return PLL_FAILURE;
  }
  if (pidx != PLLMOD_TREEINFO_PARTITION_ALL)
    ti->branch_lengths[pidx][m] = length;
  else if (ti->brlen_linkage != PLLMOD_COMMON_BRLEN_UNLINKED)
  {
    pllmod_utree_set_length(edge, length);
    ti->branch_lengths[0][m] = length;
  }
  else
    for (unsigned int p = 0; p < ti->partition_count; ++p)
      if (ti->partitions[p]) ti->branch_lengths[p][m] = length;
  return PLL_SUCCESS;
}

/* ---- topology snapshots (treeinfo.c:547-701): used by SPR rollback ---------------------------- */

extern "C" int pllmod_treeinfo_destroy_topology(pllmod_treeinfo_topology_t * topol)
{
  if (topol)
  {
    free(topol->edges);
    if (topol->branch_lengths)
    {
      for (unsigned int i = 0; i < topol->brlen_set_count; ++i) free(topol->branch_lengths[i]);
      free(topol->branch_lengths);
    }
    free(topol);
  }
  return PLL_SUCCESS;
}

extern "C" pllmod_treeinfo_topology_t * pllmod_treeinfo_get_topology(
    const pllmod_treeinfo_t * ti, pllmod_treeinfo_topology_t * topol)
{
  const unsigned int sets =
      ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED ? ti->init_partition_count : 0;
  if (!topol)
  {
    topol = (pllmod_treeinfo_topology_t *)calloc(1, sizeof(*topol));
    if (!topol)
    {
      pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for topology structure\n");
      return nullptr;
    }
    topol->edge_count = ti->tree->edge_count;
    topol->brlen_set_count = sets;
    topol->edges = (pllmod_treeinfo_edge_t *)calloc(topol->edge_count, sizeof(pllmod_treeinfo_edge_t));
    bool ok = topol->edges != nullptr;
    if (ok && sets)
    {
      topol->branch_lengths = (double **)calloc(sets, sizeof(double *));
      ok = topol->branch_lengths != nullptr;
      for (unsigned int i = 0; ok && i < sets; ++i)
        ok = (topol->branch_lengths[i] = (double *)calloc(topol->edge_count, sizeof(double))) != nullptr;
    }
    if (!ok)
    {
      pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate memory for topology buffers\n");
      pllmod_treeinfo_destroy_topology(topol);
      return nullptr;
    }
  }
  else if (ti->tree->edge_count != topol->edge_count || sets != topol->brlen_set_count)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Incompatible treeinfo_topology structure!\n");
    return nullptr;
  }
  topol->root_index = ti->root->node_index;
  unsigned int n = 0;
  for (unsigned int i = 0; i < ti->subnode_count; ++i)
  {
    const pll_unode_t * sn = ti->subnodes[i];
    if (sn->node_index < sn->back->node_index)
    {
      pllmod_treeinfo_edge_t & e = topol->edges[n++];
      e.pmatrix_index = sn->pmatrix_index;
      e.left_index = sn->node_index;
      e.right_index = sn->back->node_index;
      e.brlen = sn->length;
    }
  }
  assert(n == topol->edge_count);
  for (unsigned int i = 0; i < sets; ++i)
    memcpy(topol->branch_lengths[i], ti->branch_lengths[ti->init_partition_idx[i]],
           sizeof(double) * topol->edge_count);
  return topol;
}

extern "C" int pllmod_treeinfo_set_topology(pllmod_treeinfo_t * ti,
                                            const pllmod_treeinfo_topology_t * topol)
{
  if (!ti || !topol)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "treeinfo and/or topology is empty!\n");
    return PLL_FAILURE;
  }
  const unsigned int sets =
      ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED ? ti->init_partition_count : 0;
  if (ti->tree->edge_count != topol->edge_count || sets != topol->brlen_set_count)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Incompatible topology!\n");
    return PLL_FAILURE;
  }
  for (unsigned int i = 0; i < topol->edge_count; ++i)
  {
    const pllmod_treeinfo_edge_t & e = topol->edges[i];
    pll_unode_t * l = ti->subnodes[e.left_index];
    pll_unode_t * r = ti->subnodes[e.right_index];
    pllmod_utree_connect_nodes(l, r, e.brlen);
    l->pmatrix_index = r->pmatrix_index = e.pmatrix_index;
    if (!topol->branch_lengths) ti->branch_lengths[0][e.pmatrix_index] = e.brlen;
  }
  for (unsigned int i = 0; i < sets; ++i)
    memcpy(ti->branch_lengths[ti->init_partition_idx[i]], topol->branch_lengths[i],
           sizeof(double) * topol->edge_count);
  ti->root = ti->tree->vroot = ti->subnodes[topol->root_index];
  pllmod_treeinfo_invalidate_all(ti);
  return PLL_SUCCESS;
}

/* ---- teardown --------------------------------------------------------------------------------- */

extern "C" int pllmod_treeinfo_destroy_partition(pllmod_treeinfo_t * ti, unsigned int pidx)
{
  if (!ti)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Treeinfo structure is NULL\n");
    return PLL_FAILURE;
  }
  if (pidx >= ti->partition_count)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Partition %d is out of bounds\n", pidx);
    return PLL_FAILURE;
  }
  free(ti->clv_valid[pidx]);
  ti->clv_valid[pidx] = nullptr;
  free(ti->pmatrix_valid[pidx]);
  ti->pmatrix_valid[pidx] = nullptr;
  free(ti->param_indices[pidx]);
  ti->param_indices[pidx] = nullptr;
  free(ti->subst_matrix_symmetries[pidx]);
  ti->subst_matrix_symmetries[pidx] = nullptr;
  free(ti->deriv_precomp[pidx]);
  ti->deriv_precomp[pidx] = nullptr;
  return PLL_SUCCESS;
}

extern "C" void pllmod_treeinfo_destroy(pllmod_treeinfo_t * ti)
{
  if (!ti) return;
  free(ti->travbuffer);
  free(ti->matrix_indices);
  free(ti->operations);
  free(ti->subnodes);
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED && ti->branch_lengths) free(ti->branch_lengths[p]);
    if (ti->clv_valid && ti->pmatrix_valid && ti->param_indices && ti->subst_matrix_symmetries &&
        ti->deriv_precomp)
      pllmod_treeinfo_destroy_partition(ti, p);
  }
  free(ti->subst_matrix_symmetries);
  free(ti->constraint);
  free(ti->clv_valid);
  free(ti->pmatrix_valid);
  free(ti->linked_branch_lengths);
  free(ti->params_to_optimize);
  free(ti->alphas);
  free(ti->gamma_mode);
  free(ti->param_indices);
  free(ti->branch_lengths);
  free(ti->partition_loglh);
  free(ti->deriv_precomp);
  free(ti->brlen_scalers);
  free(ti->partitions);
  free(ti->init_partitions);
  free(ti->init_partition_idx);
  if (ti->tree)
  {
    free(ti->tree->nodes); /* the unodes belong to the caller (treeinfo.c:823-827) */
    free(ti->tree);
  }
  free(ti);
}

/* ---- validity flags --------------------------------------------------------------------------- */

extern "C" void pllmod_treeinfo_invalidate_all(pllmod_treeinfo_t * ti)
{
  const unsigned int clv_count = ti->tip_count + (ti->tip_count - 2) * 3;
  const unsigned int pm_count = ti->tree->edge_count;
  for (unsigned int i = 0; i < ti->init_partition_count; ++i)
  {
    const unsigned int p = ti->init_partition_idx[i];
    if (!part_active(ti, p)) continue;
    memset(ti->pmatrix_valid[p], 0, pm_count);
    memset(ti->clv_valid[p], 0, clv_count);
  }
}

extern "C" int pllmod_treeinfo_validate_clvs(pllmod_treeinfo_t * ti, pll_unode_t ** travbuffer,
                                             unsigned int travbuffer_size)
{
  for (unsigned int i = 0; i < ti->init_partition_count; ++i)
  {
    const unsigned int p = ti->init_partition_idx[i];
    if (!part_active(ti, p)) continue;
    for (unsigned int j = 0; j < travbuffer_size; ++j)
    {
      const pll_unode_t * node = travbuffer[j];
      if (!node->next) continue;
      /* one CLV buffer per inner node: validating one orientation invalidates the other two */
      ti->clv_valid[p][node->node_index] = 1;
      ti->clv_valid[p][node->next->node_index] = 0;
      ti->clv_valid[p][node->next->next->node_index] = 0;
    }
  }
  return PLL_SUCCESS;
}

extern "C" void pllmod_treeinfo_invalidate_pmatrix(pllmod_treeinfo_t * ti, const pll_unode_t * edge)
{
  for (unsigned int i = 0; i < ti->init_partition_count; ++i)
  {
    const unsigned int p = ti->init_partition_idx[i];
    if (ti->pmatrix_valid[p] && part_active(ti, p)) ti->pmatrix_valid[p][edge->pmatrix_index] = 0;
  }
}

extern "C" void pllmod_treeinfo_invalidate_clv(pllmod_treeinfo_t * ti, const pll_unode_t * edge)
{
  for (unsigned int i = 0; i < ti->init_partition_count; ++i)
  {
    const unsigned int p = ti->init_partition_idx[i];
    if (ti->clv_valid[p] && part_active(ti, p)) ti->clv_valid[p][edge->node_index] = 0;
  }
}

/* ---- P-matrices: batched per partition -------------------------------------------------------- */

extern "C" int pllmod_treeinfo_update_prob_matrices(pllmod_treeinfo_t * ti, int update_all)
{
  const unsigned int pm_count = ti->tree->edge_count;
  std::vector<unsigned int> idx;
  std::vector<double> len;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!part_active(ti, p) || !ti->partitions[p]) continue;
    idx.clear();
    len.clear();
    for (unsigned int m = 0; m < pm_count; ++m)
    {
      if (ti->pmatrix_valid[p][m] && !update_all) continue;
      double l = ti->branch_lengths[p][m];
      if (ti->brlen_linkage == PLLMOD_COMMON_BRLEN_SCALED) l *= ti->brlen_scalers[p];
      idx.push_back(m);
      len.push_back(l);
    }
    if (idx.empty()) continue;
    if (!pll_update_prob_matrices(ti->partitions[p], ti->param_indices[p], idx.data(), len.data(),
                                  (unsigned int)idx.size()))
      return PLL_FAILURE;
    for (unsigned int m : idx) ti->pmatrix_valid[p][m] = 1;
  }
  return PLL_SUCCESS;
}

/* ---- traversal + likelihood ------------------------------------------------------------------- */

/* shared front half of compute_loglh / update_partials: traversal, branch-length collection,
 * P-matrix refresh, op list, queued CLV updates for every local partition (treeinfo.c:966-1041) */
static int queue_traversal(pllmod_treeinfo_t * ti, int incremental, int update_pmatrices)
{
  assert(!pllmod_utree_is_tip(ti->root));
  unsigned int traversal_size = 0, ops_count = 0;
  const int collect_brlen = ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED ? 0 : 1;
  pllmod_treeinfo_set_active_partition(ti, PLLMOD_TREEINFO_PARTITION_ALL);

  if (!incremental || (update_pmatrices && collect_brlen))
    if (!pll_utree_traverse(ti->root, PLL_TREE_TRAVERSE_POSTORDER, cb_full_traversal,
                            ti->travbuffer, &traversal_size))
      return PLL_FAILURE;
  /* many partitions: P-matrices and CLV updates of all of them in one launch per kernel kind
   * (device_ctx.cu, batch mode; 0 = these partitions cannot be batched, nothing changes) */
  struct BatchScope
  {
    bool on;
    ~BatchScope() { if (on) pllb200_batch_end(); }
  } scope{false};
  if (ti->partition_count > 1)
  {
    const int b = pllb200_batch_begin(ti->partitions, ti->partition_count);
    if (b < 0) return PLL_FAILURE;
    scope.on = b == 1;
  }
  if (update_pmatrices)
  {
    if (collect_brlen)
      for (unsigned int i = 0; i < traversal_size; ++i)
        ti->branch_lengths[0][ti->travbuffer[i]->pmatrix_index] = ti->travbuffer[i]->length;
    if (!pllmod_treeinfo_update_prob_matrices(ti, !incremental)) return PLL_FAILURE;
    if (scope.on && !pllb200_batch_flush()) return PLL_FAILURE;
  }
  if (incremental)
    if (!pll_utree_traverse(ti->root, PLL_TREE_TRAVERSE_POSTORDER, cb_partial_traversal,
                            ti->travbuffer, &traversal_size))
      return PLL_FAILURE;
  pll_utree_create_operations(ti->travbuffer, traversal_size, nullptr, nullptr, ti->operations,
                              nullptr, &ops_count);
  ti->counter += ops_count;

  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    if (!ti->partitions[p]) continue;
    pllmod_treeinfo_set_active_partition(ti, (int)p);
    pll_errno = 0;
    pll_update_partials(ti->partitions[p], ti->operations, ops_count);
    if (pll_errno) return PLL_FAILURE;
    pllmod_treeinfo_validate_clvs(ti, ti->travbuffer, traversal_size);
  }
  if (scope.on)
  {
    scope.on = false;
    if (!pllb200_batch_end()) return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}

static double compute_loglh(pllmod_treeinfo_t * ti, int incremental, int update_pmatrices,
                            double ** persite_lnl)
{
  const double LOGLH_NONE = (double)NAN;
  const int old_active = ti->active_partition;
  double total = 0.0;
  if (!queue_traversal(ti, incremental, update_pmatrices))
  {
    pllmod_treeinfo_set_active_partition(ti, old_active);
    return LOGLH_NONE;
  }
  const pll_unode_t * r = ti->root;
  bool failed = false;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
  {
    ti->partition_loglh[p] = 0.0;
    if (!ti->partitions[p]) continue; /* remote partition (treeinfo.c:1024-1029) */
    if (persite_lnl && persite_lnl[p])
      ti->partition_loglh[p] = pll_compute_edge_loglikelihood(
          ti->partitions[p], r->clv_index, r->scaler_index, r->back->clv_index,
          r->back->scaler_index, r->pmatrix_index, ti->param_indices[p], persite_lnl[p]);
    else if (!pllb200_edge_loglikelihood_async(ti->partitions[p], r->clv_index, r->scaler_index,
                                               r->back->clv_index, r->back->scaler_index,
                                               r->pmatrix_index, ti->param_indices[p],
                                               &ti->partition_loglh[p]))
      failed = true;
  }
  /* one wait per partition stream, after everything has been queued */
  for (unsigned int p = 0; p < ti->partition_count; ++p)
    if (ti->partitions[p] && !pllb200_partition_wait_results(ti->partitions[p])) failed = true;
  pllmod_treeinfo_set_active_partition(ti, old_active);
  if (failed) return LOGLH_NONE;

  if (ti->parallel_reduce_cb)
    ti->parallel_reduce_cb(ti->parallel_context, ti->partition_loglh, ti->partition_count,
                           PLLMOD_COMMON_REDUCE_SUM);
  for (unsigned int p = 0; p < ti->partition_count; ++p) total += ti->partition_loglh[p];
  assert(total < 0.);
  {
    /* debugging aid (PLLB200_TRACE_LOGLH=file): every evaluation's result, bit-exact, one per line */
    static FILE * trace = getenv("PLLB200_TRACE_LOGLH") ? fopen(getenv("PLLB200_TRACE_LOGLH"), "w") : nullptr;
    if (trace) fprintf(trace, "%a %d %u\n", total, incremental, r->node_index);
  }
  return total;
}

extern "C" double pllmod_treeinfo_compute_loglh(pllmod_treeinfo_t * ti, int incremental)
{
  return compute_loglh(ti, incremental, 1, nullptr);
}

extern "C" double pllmod_treeinfo_compute_loglh_flex(pllmod_treeinfo_t * ti, int incremental,
                                                     int update_pmatrices)
{
  return compute_loglh(ti, incremental, update_pmatrices, nullptr);
}

extern "C" double pllmod_treeinfo_compute_loglh_persite(pllmod_treeinfo_t * ti, int incremental,
                                                        double ** persite_lnl)
{
  return compute_loglh(ti, incremental, 1, persite_lnl);
}

extern "C" int pllmod_treeinfo_update_partials(pllmod_treeinfo_t * ti, int incremental)
{
  const int old_active = ti->active_partition;
  const int rc = queue_traversal(ti, incremental, 1);
  pllmod_treeinfo_set_active_partition(ti, old_active);
  return rc;
}

/* ---- branch scaling (treeinfo.c:1101-1197) ---------------------------------------------------- */

extern "C" int pllmod_treeinfo_scale_branches_all(pllmod_treeinfo_t * ti, double scaler)
{
  for (unsigned int i = 0; i < ti->subnode_count; ++i) ti->subnodes[i]->length *= scaler;
  const unsigned int sets = ti->brlen_linkage == PLLMOD_COMMON_BRLEN_UNLINKED ? ti->partition_count : 1;
  for (unsigned int p = 0; p < sets; ++p)
    for (unsigned int i = 0; i < ti->tree->edge_count; ++i) ti->branch_lengths[p][i] *= scaler;
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_scale_branches_partition(pllmod_treeinfo_t * ti, unsigned int pidx,
                                                        double scaler)
{
  if (ti->brlen_linkage != PLLMOD_COMMON_BRLEN_UNLINKED)
  {
    pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE,
                     "Per-partition branch length scaling is supported in unlinked branch length "
                     "mode only.\n");
    return PLL_FAILURE;
  }
  if (pidx >= ti->partition_count)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Partition ID out of bounds\n");
    return PLL_FAILURE;
  }
  for (unsigned int i = 0; i < ti->tree->edge_count; ++i) ti->branch_lengths[pidx][i] *= scaler;
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_normalize_brlen_scalers(pllmod_treeinfo_t * ti)
{
  double sum_scalers = 0., sum_sites = 0.;
  for (unsigned int p = 0; p < ti->partition_count; ++p)
    if (ti->partitions[p])
    {
      const double w = ti->partitions[p]->pattern_weight_sum;
      sum_sites += w;
      sum_scalers += ti->brlen_scalers[p] * w;
    }
  if (ti->parallel_reduce_cb)
  {
    ti->parallel_reduce_cb(ti->parallel_context, &sum_scalers, 1, PLLMOD_COMMON_REDUCE_SUM);
    ti->parallel_reduce_cb(ti->parallel_context, &sum_sites, 1, PLLMOD_COMMON_REDUCE_SUM);
  }
  const double mean_rate = sum_scalers / sum_sites;
  pllmod_treeinfo_scale_branches_all(ti, mean_rate);
  for (unsigned int p = 0; p < ti->partition_count; ++p)
    if (ti->partitions[p]) ti->brlen_scalers[p] /= mean_rate;
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_set_tree(pllmod_treeinfo_t * ti, pll_utree_t * tree)
{
  if (!check_tree(ti, tree)) return PLL_FAILURE;
  const unsigned int node_count = tree->tip_count + tree->inner_count;
  pll_unode_t ** nodes = ti->tree->nodes; /* shallow copy: unodes are shared (treeinfo.c:1301-1311) */
  memcpy(ti->tree, tree, sizeof(pll_utree_t));
  ti->tree->nodes = nodes;
  memcpy(ti->tree->nodes, tree->nodes, sizeof(pll_unode_t *) * node_count);
  return init_tree(ti);
}


/* ---- marginal ancestral states at every inner node (treeinfo.c:1558-1712) ------------------------ */

static int anc_visit_all(pll_unode_t *) { return 1; }

/* treeinfo.c:509-544 */
extern "C" pll_utree_t * pllmod_treeinfo_get_partition_tree(const pllmod_treeinfo_t * ti, int partition_index)
{
  pll_utree_t * ptree = pll_utree_clone(ti->tree);
  if (!ptree) return nullptr;
  if (ti->brlen_linkage != PLLMOD_COMMON_BRLEN_UNLINKED) return ptree;
  const double * lengths = ti->branch_lengths[partition_index];
  const unsigned int node_count = ptree->tip_count + ptree->inner_count;
  for (unsigned int i = 0; i < node_count; ++i)
  {
    pll_unode_t * first = ptree->nodes[i];
    pll_unode_t * sub = first;
    do
    {
      if (sub->pmatrix_index > ptree->edge_count)
      {
        pllmod_set_error(PLLMOD_TREE_ERROR_INVALID_TREE,
                         "p-matrix index out of bounds (%u). treeinfo structure requires that each branch is "
                         "assigned a unique p-matrix index between 0 and branch_count-1\n", sub->pmatrix_index);
        pll_utree_destroy(ptree, nullptr);
        return nullptr;
      }
      sub->length = lengths[sub->pmatrix_index];
      sub = sub->next;
    } while (sub && sub != first);
  }
  return ptree;
}

/* ---- topological constraints: the legacy group vector (treeinfo.c:1298-1326,1440-1520) -------------- */

extern "C" int pllmod_treeinfo_set_constraint_clvmap(pllmod_treeinfo_t * ti, const int * clv_index_map)
{
  const unsigned int nodes = ti->tree->tip_count + ti->tree->inner_count;
  if (!ti->constraint)
  {
    ti->constraint = (unsigned int *)malloc(nodes * sizeof(unsigned int));
    if (!ti->constraint)
    {
      pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Can't allocate memory for topological constraint\n");
      return PLL_FAILURE;
    }
  }
  /* group id 0 = freely movable (map value -1) */
  for (unsigned int i = 0; i < nodes; ++i)
  {
    const unsigned int clv = ti->tree->nodes[i]->clv_index;
    ti->constraint[clv] = (unsigned int)(clv_index_map[clv] + 1);
  }
  return PLL_SUCCESS;
}

extern "C" int pllmod_treeinfo_set_constraint_tree(pllmod_treeinfo_t *, const pll_utree_t *, int)
{
  pllmod_set_error(PLLMOD_ERROR_NOT_IMPLEMENTED,
                   "Constraint trees need the split-set machinery (utree_constraint.c), which libpll_b200 does not "
                   "build; use pllmod_treeinfo_set_constraint_clvmap");
  return PLL_FAILURE;
}

/* group of the subtree behind `node`: its own id, else what its children agree on (s: the group being moved) */
static unsigned int constraint_group(const pll_unode_t * node, const unsigned int * cons, unsigned int s)
{
  const unsigned int own = cons[node->clv_index];
  if (!node->next || own > 0) return own;
  const unsigned int l = constraint_group(node->next->back, cons, s);
  const unsigned int r = constraint_group(node->next->next->back, cons, s);
  if (l == r) return l;
  if (!s) return l > r ? l : r;
  return (l == 0 || l == s) ? r : l;
}

static bool constraint_free(const pll_unode_t * node, const unsigned int * cons)
{
  if (!node->next) return cons[node->clv_index] == 0;
  return constraint_free(node->next->back, cons) && constraint_free(node->next->next->back, cons);
}

extern "C" int pllmod_treeinfo_constraint_check_spr(pllmod_treeinfo_t * ti, pll_unode_t * subtree,
                                                    pll_unode_t * regraft_edge)
{
  if (ti->constraint)
  {
    unsigned int s = ti->constraint[subtree->clv_index];
    if (!s) s = constraint_group(subtree->back, ti->constraint, 0);
    if (constraint_free(subtree->back, ti->constraint) || !s) return PLL_SUCCESS;
    const unsigned int r1 = constraint_group(regraft_edge, ti->constraint, s);
    const unsigned int r2 = constraint_group(regraft_edge->back, ti->constraint, s);
    return (s == r1 || s == r2) ? PLL_SUCCESS : PLL_FAILURE;
  }
  if (ti->cons_splits)
  {
    pllmod_set_error(PLLMOD_ERROR_NOT_IMPLEMENTED, "Split-based constraint checks are not built");
    return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}

/* without constraint splits the reference answers PLL_SUCCESS from all three (treeinfo.c:1522-1556) */
static int no_splits_or_fail(const pllmod_treeinfo_t * ti)
{
  if (!ti->cons_splits) return PLL_SUCCESS;
  pllmod_set_error(PLLMOD_ERROR_NOT_IMPLEMENTED, "Split-based constraint checks are not built");
  return PLL_FAILURE;
}
extern "C" int pllmod_treeinfo_constraint_subtree_affected(pllmod_treeinfo_t * ti, pll_unode_t *)
{
  return no_splits_or_fail(ti);
}
extern "C" int pllmod_treeinfo_constraint_update_splits(pllmod_treeinfo_t * ti) { return no_splits_or_fail(ti); }
extern "C" int pllmod_treeinfo_constraint_check_current(pllmod_treeinfo_t * ti) { return no_splits_or_fail(ti); }

extern "C" void pllmod_treeinfo_destroy_ancestral(pllmod_ancestral_t * ancestral)
{
  if (!ancestral) return;
  free(ancestral->nodes);
  free(ancestral->partition_indices);
  if (ancestral->probs)
    for (unsigned int i = 0; i < ancestral->node_count; ++i) free(ancestral->probs[i]);
  free(ancestral->probs);
  if (ancestral->tree) pll_utree_destroy(ancestral->tree, nullptr);
  free(ancestral);
}

extern "C" pllmod_ancestral_t * pllmod_treeinfo_compute_ancestral(pllmod_treeinfo_t * ti)
{
  const unsigned int inner = ti->tree->inner_count, total = ti->tree->tip_count + inner;
  pllmod_ancestral_t * anc = (pllmod_ancestral_t *)calloc(1, sizeof(pllmod_ancestral_t));
  if (!anc)
  {
    pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Can't allocate memory for ancestral probabilities\n");
    return nullptr;
  }
  size_t span = 0; /* doubles per node: all (local) partitions back to back, [site][state] each */
  for (unsigned int p = 0; p < ti->init_partition_count; ++p)
    span += (size_t)ti->init_partitions[p]->sites * ti->init_partitions[p]->states;
  anc->node_count = inner;
  anc->partition_count = ti->init_partition_count;
  anc->nodes = (pll_unode_t **)calloc(inner, sizeof(pll_unode_t *));
  anc->partition_indices = (unsigned int *)calloc(anc->partition_count ? anc->partition_count : 1, sizeof(unsigned int));
  anc->probs = (double **)calloc(inner, sizeof(double *));
  anc->tree = pll_utree_clone(ti->tree);
  bool ok = anc->nodes && anc->partition_indices && anc->probs && anc->tree;
  for (unsigned int i = 0; ok && i < inner; ++i) ok = (anc->probs[i] = (double *)calloc(span ? span : 1, sizeof(double))) != nullptr;
  std::vector<pll_unode_t *> order(total);
  unsigned int visited = 0;
  ok = ok && pll_utree_traverse(anc->tree->vroot, PLL_TREE_TRAVERSE_POSTORDER, anc_visit_all, order.data(), &visited) &&
       visited == total;
  if (!ok)
  {
    pllmod_treeinfo_destroy_ancestral(anc);
    if (!pll_errno) pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Can't allocate memory for ancestral probabilities\n");
    return nullptr;
  }
  /* inner nodes of the cloned tree in post-order; unnamed ones are called Node1, Node2, ... */
  unsigned int n = 0;
  for (pll_unode_t * node : order)
  {
    if (!node->next) continue;
    if (!node->label)
    {
      char name[100];
      snprintf(name, sizeof(name), "Node%u", n + 1);
      node->label = node->next->label = node->next->next->label = strdup(name);
    }
    anc->nodes[n++] = node;
  }

  pll_unode_t * old_root = ti->root;
  for (unsigned int i = 0; i < inner; ++i)
  {
    const pll_unode_t * node = anc->nodes[i];
    /* orient all CLVs towards this node (partial traversal), then combine the two ends of its edge */
    ti->root = ti->subnodes[node->node_index];
    pllmod_treeinfo_compute_loglh(ti, 1);
    double * out = anc->probs[i];
    for (unsigned int p = 0; p < ti->init_partition_count; ++p)
    {
      pll_partition_t * part = ti->init_partitions[p];
      const unsigned int pidx = ti->init_partition_idx[p];
      anc->partition_indices[p] = pidx;
      if (!pll_compute_node_ancestral(part, node->clv_index, node->scaler_index, node->back->clv_index,
                                      node->back->scaler_index, node->pmatrix_index, ti->param_indices[pidx], out))
      {
        ti->root = old_root;
        pllmod_treeinfo_destroy_ancestral(anc);
        return nullptr;
      }
      out += (size_t)part->sites * part->states;
    }
  }
  ti->root = old_root;
  return anc;
}
