/*
 * utree.cpp — the small slice of unrooted-tree plumbing the treeinfo / optimiser drivers need.
 * Host pointer surgery only (no likelihood arithmetic).  Reference behaviour:
 *   pll_utree_wraptree          libs/libpll/src/parse_utree.y:436-535 (node order: tips first, then
 *                               inner nodes, both in a depth-first walk from the root)
 *   pll_utree_destroy           parse_utree.y:107-145
 *   pll_utree_graph_destroy     parse_utree.y:68-105
 *   pll_utree_graph_clone/clone utree.c:549-633
 *   pll_utree_traverse          libs/libpll/src/utree.c:394-459 (root->back subtree first, then root)
 *   pll_utree_create_operations utree.c:317-366
 *   pllmod_utree_create_random  src/tree/pll_tree.c:903-980 (random branch insertion)
 *   pllmod_utree_connect_nodes / set_length / is_tip   src/tree/utree_operations.c
 */
#include "internal.h"

#include <cstdlib>
#include <cstring>
#include <vector>

extern "C" int pllmod_utree_is_tip(const pll_unode_t * node) { return node->next == nullptr; }

extern "C" void pllmod_utree_set_length(pll_unode_t * edge, double length)
{
  edge->length = edge->back->length = length;
}

extern "C" int pllmod_utree_connect_nodes(pll_unode_t * parent, pll_unode_t * child, double length)
{
  if (!parent || !child) return PLL_FAILURE;
  parent->back = child;
  child->back = parent;
  pllmod_utree_set_length(parent, length);
  /* the edge takes the p-matrix index of the parent end (utree_operations.c:370-371) */
  child->pmatrix_index = parent->pmatrix_index;
  return PLL_SUCCESS;
}

static void collect_nodes(pll_unode_t * node, bool is_root, std::vector<pll_unode_t *> & tips,
                          std::vector<pll_unode_t *> & inner)
{
  if (!node->next)
  {
    tips.push_back(node);
    return;
  }
  pll_unode_t * sn = is_root ? node : node->next;
  do
  {
    collect_nodes(sn->back, false, tips, inner);
    sn = sn->next;
  } while (sn != node);
  inner.push_back(node);
}

extern "C" pll_utree_t * pll_utree_wraptree(pll_unode_t * root, unsigned int tip_count)
{
  if (!root)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "NULL tree root.");
    return nullptr;
  }
  if (tip_count < 3 && tip_count != 0)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid tip_count value (%u).", tip_count);
    return nullptr;
  }
  if (!root->next) root = root->back;
  std::vector<pll_unode_t *> tips, inner;
  collect_nodes(root, true, tips, inner);
  if (tip_count && tips.size() != tip_count)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Tree has %zu tips, expected %u.", tips.size(), tip_count);
    return nullptr;
  }
  if (inner.size() + 2 != tips.size())
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Input tree is not strictly bifurcating.");
    return nullptr;
  }
  pll_utree_t * tree = (pll_utree_t *)malloc(sizeof(pll_utree_t));
  pll_unode_t ** nodes = (pll_unode_t **)malloc(sizeof(pll_unode_t *) * (tips.size() + inner.size()));
  if (!tree || !nodes)
  {
    free(tree);
    free(nodes);
    pllb200_set_error(PLL_ERROR_MEM_ALLOC, "Unable to allocate enough memory.");
    return nullptr;
  }
  size_t q = 0;
  for (pll_unode_t * n : tips) nodes[q++] = n;
  for (pll_unode_t * n : inner) nodes[q++] = n;
  tree->tip_count = (unsigned int)tips.size();
  tree->inner_count = (unsigned int)inner.size();
  tree->edge_count = tree->tip_count + tree->inner_count - 1;
  tree->binary = 1;
  tree->nodes = nodes;
  tree->vroot = root;
  return tree;
}

extern "C" void pll_utree_destroy(pll_utree_t * tree, void (*cb_destroy)(void *))
{
  if (!tree) return;
  for (unsigned int i = 0; i < tree->tip_count + tree->inner_count; ++i)
  {
    pll_unode_t * first = tree->nodes[i];
    if (!first) continue;
    free(first->label);
    pll_unode_t * n = first;
    do
    {
      pll_unode_t * nx = n->next;
      if (cb_destroy && n->data) cb_destroy(n->data);
      free(n);
      n = nx;
    } while (n && n != first);
  }
  free(tree->nodes);
  free(tree);
}

/* Free a node graph that has no pll_utree_t wrapper (parse_utree.y:68-105): everything reachable
 * from `root`, both sides of its edge.  The records of an inner node share one label. */
static void free_subtree(pll_unode_t * entry, bool also_behind_entry, void (*cb_destroy)(void *))
{
  if (!entry->next)
  {
    if (cb_destroy && entry->data) cb_destroy(entry->data);
    free(entry->label);
    free(entry);
    return;
  }
  free(entry->label);
  pll_unode_t * n = entry;
  do
  {
    pll_unode_t * nx = n->next;
    if ((n != entry || also_behind_entry) && n->back) free_subtree(n->back, false, cb_destroy);
    if (cb_destroy && n->data) cb_destroy(n->data);
    free(n);
    n = nx;
  } while (n && n != entry);
}

extern "C" void pll_utree_graph_destroy(pll_unode_t * root, void (*cb_destroy)(void *))
{
  if (root) free_subtree(root, true, cb_destroy);
}

/* Deep copy of a node graph (utree.c:549-633): every ring is duplicated with ONE copy of its label,
 * `data` pointers are carried over as they are. */
static pll_unode_t * clone_ring(const pll_unode_t * node)
{
  pll_unode_t * first = nullptr;
  pll_unode_t * prev = nullptr;
  char * label = node->label ? strdup(node->label) : nullptr;
  const pll_unode_t * n = node;
  do
  {
    pll_unode_t * copy = (pll_unode_t *)malloc(sizeof(pll_unode_t));
    *copy = *n;
    copy->label = label;
    copy->back = nullptr;
    copy->next = nullptr;
    if (prev)
      prev->next = copy;
    else
      first = copy;
    prev = copy;
    n = n->next;
  } while (n && n != node);
  if (node->next) prev->next = first;
  return first;
}

static void clone_behind(pll_unode_t * copy, const pll_unode_t * orig)
{
  if (!orig->back) return;
  pll_unode_t * far = clone_ring(orig->back);
  copy->back = far;
  far->back = copy;
  const pll_unode_t * o = orig->back->next;
  pll_unode_t * c = far->next;
  while (o && o != orig->back)
  {
    clone_behind(c, o);
    o = o->next;
    c = c->next;
  }
}

extern "C" pll_unode_t * pll_utree_graph_clone(const pll_unode_t * root)
{
  pll_unode_t * copy = clone_ring(root);
  const pll_unode_t * o = root;
  pll_unode_t * c = copy;
  do
  {
    clone_behind(c, o);
    o = o->next;
    c = c->next;
  } while (o && o != root);
  return copy;
}

extern "C" pll_utree_t * pll_utree_clone(const pll_utree_t * tree)
{
  return pll_utree_wraptree(pll_utree_graph_clone(tree->vroot), tree->tip_count);
}

static void traverse_rec(pll_unode_t * node, int traversal, int (*cb)(pll_unode_t *),
                         unsigned int * index, pll_unode_t ** out)
{
  if (!cb(node)) return;
  if (traversal == PLL_TREE_TRAVERSE_PREORDER) out[(*index)++] = node;
  if (node->next)
  {
    pll_unode_t * sn = node->next;
    do
    {
      traverse_rec(sn->back, traversal, cb, index, out);
      sn = sn->next;
    } while (sn && sn != node);
  }
  if (traversal == PLL_TREE_TRAVERSE_POSTORDER) out[(*index)++] = node;
}

extern "C" int pll_utree_traverse(pll_unode_t * root, int traversal, int (*cbtrav)(pll_unode_t *),
                                  pll_unode_t ** outbuffer, unsigned int * trav_size)
{
  *trav_size = 0;
  if (!root->next) return PLL_FAILURE;
  if (traversal != PLL_TREE_TRAVERSE_POSTORDER && traversal != PLL_TREE_TRAVERSE_PREORDER)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "Invalid traversal value.");
    return PLL_FAILURE;
  }
  traverse_rec(root->back, traversal, cbtrav, trav_size, outbuffer);
  traverse_rec(root, traversal, cbtrav, trav_size, outbuffer);
  return PLL_SUCCESS;
}

extern "C" void pll_utree_create_operations(pll_unode_t * const * trav_buffer,
                                            unsigned int trav_buffer_size, double * branches,
                                            unsigned int * pmatrix_indices, pll_operation_t * ops,
                                            unsigned int * matrix_count, unsigned int * ops_count)
{
  *ops_count = 0;
  if (matrix_count) *matrix_count = 0;
  if (!trav_buffer_size) return;
  const pll_unode_t * root_back = trav_buffer[trav_buffer_size - 1]->back;
  for (unsigned int i = 0; i < trav_buffer_size; ++i)
  {
    const pll_unode_t * node = trav_buffer[i];
    if (node != root_back) /* the root edge is listed once, by the root itself */
    {
      if (branches) *branches++ = node->length;
      if (pmatrix_indices) *pmatrix_indices++ = node->pmatrix_index;
      if (matrix_count) ++*matrix_count;
    }
    if (node->next)
    {
      const pll_unode_t * c1 = node->next->back;
      const pll_unode_t * c2 = node->next->next->back;
      pll_operation_t & op = ops[(*ops_count)++];
      op.parent_clv_index = node->clv_index;
      op.parent_scaler_index = node->scaler_index;
      op.child1_clv_index = c1->clv_index;
      op.child1_scaler_index = c1->scaler_index;
      op.child1_matrix_index = c1->pmatrix_index;
      op.child2_clv_index = c2->clv_index;
      op.child2_scaler_index = c2->scaler_index;
      op.child2_matrix_index = c2->pmatrix_index;
    }
  }
}

/* ---- random topology ------------------------------------------------------------------------ */

static pll_unode_t * new_inner_triplet(unsigned int clv, int scaler, unsigned int * node_id)
{
  pll_unode_t * a = (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
  pll_unode_t * b = (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
  pll_unode_t * c = (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
  a->next = b;
  b->next = c;
  c->next = a;
  pll_unode_t * t[3] = {a, b, c};
  for (pll_unode_t * n : t)
  {
    n->clv_index = clv;
    n->scaler_index = scaler;
    n->pmatrix_index = clv;
    n->node_index = (*node_id)++;
  }
  return a;
}

extern "C" pll_utree_t * pllmod_utree_create_random(unsigned int taxa_count,
                                                    const char * const * names,
                                                    unsigned int random_seed)
{
  if (taxa_count < 3)
  {
    pllb200_set_error(PLL_ERROR_PARAM_INVALID, "A tree needs at least 3 taxa.");
    return nullptr;
  }
  const unsigned int inner_count = taxa_count - 2;
  std::vector<pll_unode_t *> nodes(taxa_count + inner_count);
  unsigned int node_id = 0;
  for (unsigned int i = 0; i < taxa_count; ++i)
  {
    pll_unode_t * t = (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
    t->clv_index = t->pmatrix_index = i;
    t->scaler_index = PLL_SCALE_BUFFER_NONE;
    t->node_index = node_id++;
    if (names && names[i])
    {
      t->label = (char *)malloc(strlen(names[i]) + 1);
      strcpy(t->label, names[i]);
    }
    nodes[i] = t;
  }
  for (unsigned int i = 0; i < inner_count; ++i)
    nodes[taxa_count + i] = new_inner_triplet(taxa_count + i, (int)i, &node_id);

  /* star of the first three taxa, then insert each further tip into a uniformly chosen branch */
  std::vector<pll_unode_t *> branches; /* one end per branch */
  pll_unode_t * c0 = nodes[taxa_count];
  pllmod_utree_connect_nodes(nodes[0], c0, PLLMOD_TREE_DEFAULT_BRANCH_LENGTH);
  pllmod_utree_connect_nodes(nodes[1], c0->next, PLLMOD_TREE_DEFAULT_BRANCH_LENGTH);
  pllmod_utree_connect_nodes(nodes[2], c0->next->next, PLLMOD_TREE_DEFAULT_BRANCH_LENGTH);
  branches.push_back(nodes[0]);
  branches.push_back(nodes[1]);
  branches.push_back(nodes[2]);

  unsigned long long rng = 0x9E3779B97F4A7C15ull ^ ((unsigned long long)random_seed * 0xD1B54A32D192ED03ull);
  auto next_rand = [&rng]() {
    rng ^= rng << 13;
    rng ^= rng >> 7;
    rng ^= rng << 17;
    return rng;
  };
  for (unsigned int t = 3; t < taxa_count; ++t)
  {
    pll_unode_t * inner = nodes[taxa_count + t - 2];
    pll_unode_t * tip = nodes[t];
    pll_unode_t * e = branches[next_rand() % branches.size()];
    pll_unode_t * eb = e->back;
    /* split e--eb with `inner`:  e--inner, inner.next--eb, inner.next.next--tip */
    pllmod_utree_connect_nodes(e, inner, PLLMOD_TREE_DEFAULT_BRANCH_LENGTH);
    pllmod_utree_connect_nodes(eb, inner->next, PLLMOD_TREE_DEFAULT_BRANCH_LENGTH);
    pllmod_utree_connect_nodes(tip, inner->next->next, PLLMOD_TREE_DEFAULT_BRANCH_LENGTH);
    branches.push_back(inner->next);
    branches.push_back(tip);
  }
  /* one unique p-matrix index per branch, in [0, 2n-3) */
  for (size_t i = 0; i < branches.size(); ++i)
    branches[i]->pmatrix_index = branches[i]->back->pmatrix_index = (unsigned int)i;
  return pll_utree_wraptree(c0, taxa_count);
}
