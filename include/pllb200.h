/*
 * pllb200.h — C ABI of the B200-native phylogenetic-likelihood library (libpll_b200.so).
 *
 * This header is the drop-in boundary for ONE hot path of ddarriba/pll-modules + libpll-2:
 * P-matrix construction, CLV ("partials") updates over operation lists, edge/root log-likelihood,
 * sumtable + first/second derivatives, and the pllmod_treeinfo_* / pllmod_opt_* host drivers on top.
 *
 * Every symbol below that starts with pll_ / pllmod_ keeps the NAME, ARGUMENT MEANING, RETURN
 * CONVENTION and STRUCT LAYOUT of the reference declaration it replaces, cited as (file:line)
 * relative to the reference tree:
 *     L = libs/libpll/src/pll.h      T = src/tree/pll_tree.h
 *     O = src/optimize/pll_optimize.h   C = src/pllmod_common.h
 * so that a caller compiled against the reference headers can be re-linked against
 * libpll_b200.so unchanged for this path.  Symbols that start with pllb200_ are additive.
 *
 * Residency contract (differs from the CPU library):
 *   - partition->clv[i] (allocated ones), partition->scale_buffer[i] and partition->pmatrix[i] are
 *     DEVICE pointers into HBM arenas.  The host must not dereference them; use
 *     pllb200_download_clv / _scaler / _pmatrix.
 *   - tipchars, pattern_weights, invariant, model arrays (rates, weights, frequencies, subst_params,
 *     eigen*, prop_invar) are ordinary host memory exactly as in the reference; the library mirrors
 *     them to the device when an entry point consumes them.
 *   - sumtable buffers passed to pll_update_sumtable / pll_compute_likelihood_derivatives are used
 *     as KEYS for device-resident shadow buffers; their host bytes are never touched.
 *
 * Plain C, no C++ and no CUDA/torch types in any signature.
 */
#ifndef PLLB200_H_
#define PLLB200_H_

#include <stddef.h>
#include <stdio.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PLL_EXPORT

/* ---- return codes / constants (L:82-110) ------------------------------------------------- */
#define PLL_FAILURE 0
#define PLL_SUCCESS 1
#define PLL_FALSE 0
#define PLL_TRUE 1

#define PLL_ALIGNMENT_CPU 8
#define PLL_ALIGNMENT_SSE 16
#define PLL_ALIGNMENT_AVX 32

#define PLL_ASCII_SIZE 256

/* 2^256 and 2^-256: the underflow rescaling constants (L:96-100) */
#define PLL_SCALE_FACTOR 115792089237316195423570985008687907853269984665640564039457584007913129639936.0
#define PLL_SCALE_THRESHOLD (1.0 / PLL_SCALE_FACTOR)
#define PLL_SCALE_BUFFER_NONE -1
#define PLL_SCALE_RATE_MAXDIFF 4 /* L:104 */

#define PLL_MISC_EPSILON 1e-8
#define PLL_EIGEN_MINFREQ 1e-6

/* ---- attribute flags (L:114-137).  ARCH bits are accepted and ignored: the arithmetic always
 *      runs on the device. ------------------------------------------------------------------ */
#define PLL_ATTRIB_ARCH_CPU 0
#define PLL_ATTRIB_ARCH_SSE (1 << 0)
#define PLL_ATTRIB_ARCH_AVX (1 << 1)
#define PLL_ATTRIB_ARCH_AVX2 (1 << 2)
#define PLL_ATTRIB_ARCH_AVX512 (1 << 3)
#define PLL_ATTRIB_ARCH_MASK 0xF
#define PLL_ATTRIB_PATTERN_TIP (1 << 4)
#define PLL_ATTRIB_AB_LEWIS (1 << 5)
#define PLL_ATTRIB_AB_FELSENSTEIN (2 << 5)
#define PLL_ATTRIB_AB_STAMATAKIS (3 << 5)
#define PLL_ATTRIB_AB_MASK (7 << 5)
#define PLL_ATTRIB_AB_FLAG (1 << 8)
#define PLL_ATTRIB_RATE_SCALERS (1 << 9)
#define PLL_ATTRIB_SITE_REPEATS (1 << 10) /* rejected: PLL_ERROR_PARAM_INVALID */
#define PLL_ATTRIB_MASK ((1 << 11) - 1)

#define PLL_TREE_TRAVERSE_POSTORDER 1
#define PLL_TREE_TRAVERSE_PREORDER 2

#define PLL_GAMMA_RATES_MEAN 0
#define PLL_GAMMA_RATES_MEDIAN 1

/* ---- error codes (L:162-196) -------------------------------------------------------------- */
#define PLL_ERROR_FILE_OPEN 100
#define PLL_ERROR_FILE_SEEK 101
#define PLL_ERROR_FILE_EOF 102
#define PLL_ERROR_MEM_ALLOC 112
#define PLL_ERROR_PARAM_INVALID 113
#define PLL_ERROR_TIPDATA_ILLEGALSTATE 114
#define PLL_ERROR_TIPDATA_ILLEGALFUNCTION 115
#define PLL_ERROR_INVAR_INCOMPAT 117
#define PLL_ERROR_INVAR_PROPORTION 118
#define PLL_ERROR_INVAR_PARAMINDEX 119
#define PLL_ERROR_INVAR_NONEFOUND 120
#define PLL_ERROR_AB_INVALIDMETHOD 121
#define PLL_ERROR_AB_NOSUPPORT 122
#define PLL_ERROR_EINVAL 130
#define PLL_ERROR_TREE_INVALID 133
/* additive: a CUDA runtime / device failure, message carries cudaGetErrorString */
#define PLLB200_ERROR_DEVICE 9001

typedef unsigned long long pll_state_t;
typedef int pll_bool_t;

struct pll_repeats; /* site repeats are out of scope; the pointer slot is kept for layout */

/* L:241-288 — layout identical (35 fields, same order and types). */
typedef struct pll_partition
{
  unsigned int tips;
  unsigned int clv_buffers;
  unsigned int nodes;
  unsigned int states;
  unsigned int sites;
  unsigned int pattern_weight_sum;
  unsigned int rate_matrices;
  unsigned int prob_matrices;
  unsigned int rate_cats;
  unsigned int scale_buffers;
  unsigned int attributes;

  size_t alignment;
  unsigned int states_padded;

  double ** clv;               /* DEVICE pointers (NULL for tips under PATTERN_TIP) */
  double ** pmatrix;           /* DEVICE pointers */
  double * rates;
  double * rate_weights;
  double ** subst_params;
  unsigned int ** scale_buffer; /* DEVICE pointers */
  double ** frequencies;
  double * prop_invar;
  int * invariant;
  unsigned int * pattern_weights;

  int * eigen_decomp_valid;
  double ** eigenvecs;
  double ** inv_eigenvecs;
  double ** eigenvals;

  unsigned int maxstates;
  unsigned char ** tipchars;
  unsigned char * charmap;
  double * ttlookup;           /* unused on the device path (always NULL) */
  pll_state_t * tipmap;

  int asc_bias_alloc;
  int asc_additional_sites;

  struct pll_repeats * repeats; /* always NULL */
} pll_partition_t;

/* L:325-335 */
typedef struct pll_operation
{
  unsigned int parent_clv_index;
  int parent_scaler_index;
  unsigned int child1_clv_index;
  unsigned int child1_matrix_index;
  int child1_scaler_index;
  unsigned int child2_clv_index;
  unsigned int child2_matrix_index;
  int child2_scaler_index;
} pll_operation_t;

/* L:388-400 */
typedef struct pll_unode_s
{
  char * label;
  double length;
  unsigned int node_index;
  unsigned int clv_index;
  int scaler_index;
  unsigned int pmatrix_index;
  struct pll_unode_s * next;
  struct pll_unode_s * back;
  void * data;
} pll_unode_t;

/* L:402-411 */
typedef struct pll_utree_s
{
  unsigned int tip_count;
  unsigned int inner_count;
  unsigned int edge_count;
  int binary;
  pll_unode_t ** nodes;
  pll_unode_t * vroot;
} pll_utree_t;

/* L:553-554 */
extern __thread int pll_errno;
extern __thread char pll_errmsg[200];

/* L:556-559 — state maps for tip encoding (bitmask per ASCII character) */
extern const pll_state_t pll_map_bin[256];
extern const pll_state_t pll_map_nt[256];
extern const pll_state_t pll_map_aa[256];

/* ---- libpll partition API --------------------------------------------------------------- */

/* L:638-646 (pll.c:424). Allocates the HBM arenas. Fails (NULL, pll_errno=PLLB200_ERROR_DEVICE)
 * when no CUDA device is usable: there is no CPU fallback. */
pll_partition_t * pll_partition_create(unsigned int tips,
                                       unsigned int clv_buffers,
                                       unsigned int states,
                                       unsigned int sites,
                                       unsigned int rate_matrices,
                                       unsigned int prob_matrices,
                                       unsigned int rate_cats,
                                       unsigned int scale_buffers,
                                       unsigned int attributes);
/* L:648 (pll.c:870) */
void pll_partition_destroy(pll_partition_t * partition);
/* L:650-653 (pll.c:1026) */
int pll_set_tip_states(pll_partition_t * partition, unsigned int tip_index,
                       const pll_state_t * map, const char * sequence);
/* L:655-658 (pll.c:1066) */
int pll_set_tip_clv(pll_partition_t * partition, unsigned int tip_index,
                    const double * clv, int padding);
/* L:660-661 (pll.c:1131) */
void pll_set_pattern_weights(pll_partition_t * partition, const unsigned int * pattern_weights);
/* L:674-675 (pll.c:133,148) */
void * pll_aligned_alloc(size_t size, size_t alignment);
void pll_aligned_free(void * ptr);
/* L:655-659 (pll.c:1145,1193): ascertainment-bias correction type (PLL_ATTRIB_AB_LEWIS /
 * _FELSENSTEIN / _STAMATAKIS; the partition must have been created with an AB attribute) and the
 * weights of the `states` invariant dummy sites */
int pll_set_asc_bias_type(pll_partition_t * partition, int asc_bias_type);
void pll_set_asc_state_weights(pll_partition_t * partition, const unsigned int * state_weights);
/* pll.c:1202 */
void pll_fill_parent_scaler(unsigned int scaler_size, unsigned int * parent_scaler,
                            const unsigned int * left_scaler, const unsigned int * right_scaler);

/* L:746-748 (models.c:546) */
void pll_set_subst_params(pll_partition_t * partition, unsigned int params_index,
                          const double * params);
/* L:750-752 (models.c:445) */
void pll_set_frequencies(pll_partition_t * partition, unsigned int params_index,
                         const double * frequencies);
/* L:754-755 (models.c:478) */
void pll_set_category_rates(pll_partition_t * partition, const double * rates);
/* L:757-758 (models.c:485) */
void pll_set_category_weights(pll_partition_t * partition, const double * rate_weights);
/* L:760-761 (models.c:293) host-side symmetric eigensolver */
int pll_update_eigen(pll_partition_t * partition, unsigned int params_index);
/* L:763-767 (models.c:412) */
int pll_update_prob_matrices(pll_partition_t * partition,
                             const unsigned int * params_indices,
                             const unsigned int * matrix_indices,
                             const double * branch_lengths,
                             unsigned int count);
/* L:769-770 (models.c:651) */
unsigned int pll_count_invariant_sites(pll_partition_t * partition, unsigned int * state_inv_count);
/* L:772 (models.c:651..) */
int pll_update_invariant_sites(pll_partition_t * partition);
/* L:774-776 (models.c:495) */
int pll_update_invariant_sites_proportion(pll_partition_t * partition,
                                          unsigned int params_index, double prop_invar);

/* L:784-787 (likelihood.c:122) */
double pll_compute_root_loglikelihood(pll_partition_t * partition, unsigned int clv_index,
                                      int scaler_index, const unsigned int * freqs_indices,
                                      double * persite_lnl);
/* L:789-797 (likelihood.c:586) */
double pll_compute_edge_loglikelihood(pll_partition_t * partition,
                                      unsigned int parent_clv_index, int parent_scaler_index,
                                      unsigned int child_clv_index, int child_scaler_index,
                                      unsigned int matrix_index,
                                      const unsigned int * freqs_indices,
                                      double * persite_lnl);
/* L:799-818 (likelihood.c:639-823): marginal ancestral state probabilities of an inner node,
 * ancestral[site][state], normalised per site; the _extbuf scratch arguments are ignored */
int pll_compute_node_ancestral(pll_partition_t * partition, unsigned int node_clv_index,
                               int node_scaler_index, unsigned int other_clv_index, int other_scaler_index,
                               unsigned int matrix_index, const unsigned int * freqs_indices,
                               double * ancestral);
int pll_compute_node_ancestral_extbuf(pll_partition_t * partition, unsigned int node_clv_index,
                                      int node_scaler_index, unsigned int other_clv_index,
                                      int other_scaler_index, unsigned int pmatrix_index,
                                      const unsigned int * freqs_indices, double * ancestral,
                                      double * temp_clv, unsigned int * temp_scaler, double * ident_pmat);
/* L:823-825 (partials.c:237) — asynchronous on the partition's stream */
void pll_update_partials(pll_partition_t * partition, const pll_operation_t * operations,
                         unsigned int count);
/* L:834-840 (derivatives.c:239) */
int pll_update_sumtable(pll_partition_t * partition,
                        unsigned int parent_clv_index, unsigned int child_clv_index,
                        int parent_scaler_index, int child_scaler_index,
                        const unsigned int * params_indices, double * sumtable);
/* L:842-849 (derivatives.c:333) */
int pll_compute_likelihood_derivatives(pll_partition_t * partition,
                                       int parent_scaler_index, int child_scaler_index,
                                       double branch_length,
                                       const unsigned int * params_indices,
                                       const double * sumtable,
                                       double * d_f, double * dd_f);
/* L:856-858 (gamma.c:220) */
int pll_compute_gamma_cats(double alpha, unsigned int categories, double * output_rates,
                           int rates_mode);

/* ---- libpll utree subset the treeinfo/optimize drivers need -------------------------------- */
/* L:1010-1011 (parse_utree.y:99-145) */
void pll_utree_destroy(pll_utree_t * tree, void (*cb_destroy)(void *));
/* L:927 (parse_utree.y:99) */
void pll_utree_graph_destroy(pll_unode_t * root, void (*cb_destroy)(void *));
/* L:943-945 (utree.c:605-633) */
pll_unode_t * pll_utree_graph_clone(const pll_unode_t * root);
pll_utree_t * pll_utree_clone(const pll_utree_t * tree);
/* L:1020-1021 (parse_utree.y:522) */
pll_utree_t * pll_utree_wraptree(pll_unode_t * root, unsigned int tip_count);
/* L:929-933 (utree.c:427) */
int pll_utree_traverse(pll_unode_t * root, int traversal, int (*cbtrav)(pll_unode_t *),
                       pll_unode_t ** outbuffer, unsigned int * trav_size);
/* L:921-927 (utree.c:317) */
void pll_utree_create_operations(pll_unode_t * const * trav_buffer, unsigned int trav_buffer_size,
                                 double * branches, unsigned int * pmatrix_indices,
                                 pll_operation_t * ops, unsigned int * matrix_count,
                                 unsigned int * ops_count);

/* ---- pll-modules common (C:24-47) --------------------------------------------------------- */
#define PLLMOD_COMMON_BRLEN_LINKED 0
#define PLLMOD_COMMON_BRLEN_SCALED 1
#define PLLMOD_COMMON_BRLEN_UNLINKED 2
#define PLLMOD_COMMON_REDUCE_SUM 0
#define PLLMOD_COMMON_REDUCE_MAX 1
#define PLLMOD_COMMON_REDUCE_MIN 2
#define PLLMOD_ERROR_INVALID_RANGE 1001
#define PLLMOD_ERROR_INVALID_NODE_TYPE 1002
#define PLLMOD_ERROR_INVALID_INDEX 1003
#define PLLMOD_ERROR_NOT_IMPLEMENTED 1004
void pllmod_set_error(int errno_, const char * errmsg_fmt, ...);
void pllmod_reset_error(void);

/* ---- pll-modules tree subset (T:35,63,219-277) -------------------------------------------- */
#define PLLMOD_TREE_DEFAULT_BRANCH_LENGTH 0.1
#define PLLMOD_TREE_ERROR_INVALID_TREE_SIZE 3456
#define PLLMOD_TREE_ERROR_INVALID_TREE 3584
#define PLLMOD_TREEINFO_PARTITION_ALL -1

typedef struct pllb200_split_set_opaque pll_split_set_t; /* constraint checks: out of scope */

/* T:202-208 */
typedef struct treeinfo_edge
{
  unsigned int left_index;
  unsigned int right_index;
  unsigned int pmatrix_index;
  double brlen;
} pllmod_treeinfo_edge_t;

/* T:210-217 */
typedef struct treeinfo_topology
{
  unsigned int edge_count;
  unsigned int brlen_set_count;
  unsigned int root_index;
  pllmod_treeinfo_edge_t * edges;
  double ** branch_lengths;
} pllmod_treeinfo_topology_t;

/* T:219-277 — layout identical. */
typedef struct treeinfo
{
  unsigned int tip_count;
  unsigned int partition_count;
  int brlen_linkage;
  double * linked_branch_lengths;
  pll_unode_t * root;
  pll_utree_t * tree;
  unsigned int subnode_count;
  pll_unode_t ** subnodes;
  pll_partition_t ** partitions;
  double * alphas;
  int * gamma_mode;
  unsigned int ** param_indices;
  int ** subst_matrix_symmetries;
  double ** branch_lengths;
  double * brlen_scalers;
  double * partition_loglh;
  int * params_to_optimize;
  unsigned int init_partition_count;
  unsigned int * init_partition_idx;
  pll_partition_t ** init_partitions;
  unsigned int * constraint;
  pll_split_set_t * cons_splits;
  pll_split_set_t * tree_splits;
  double ** deriv_precomp;
  char ** clv_valid;
  char ** pmatrix_valid;
  pll_unode_t ** travbuffer;
  unsigned int * matrix_indices;
  pll_operation_t * operations;
  int active_partition;
  unsigned int counter;
  void * parallel_context;
  void (*parallel_reduce_cb)(void *, double *, size_t, int);
} pllmod_treeinfo_t;

/* pll_tree.c:903 — random topology, default branch lengths (RNG differs from the reference's
 * glibc-style generator: topologies are valid but not identical for equal seeds) */
pll_utree_t * pllmod_utree_create_random(unsigned int taxa_count, const char * const * names,
                                         unsigned int random_seed);
/* utree_operations.c */
int pllmod_utree_is_tip(const pll_unode_t * node);
void pllmod_utree_set_length(pll_unode_t * edge, double length);
int pllmod_utree_connect_nodes(pll_unode_t * parent, pll_unode_t * child, double length);

/* T:616-790 (treeinfo.c) */
pllmod_treeinfo_t * pllmod_treeinfo_create(pll_unode_t * root, unsigned int tips,
                                           unsigned int partitions, int brlen_linkage);
int pllmod_treeinfo_set_parallel_context(pllmod_treeinfo_t * treeinfo, void * parallel_context,
                                         void (*parallel_reduce_cb)(void *, double *, size_t, int op));
int pllmod_treeinfo_init_partition(pllmod_treeinfo_t * treeinfo, unsigned int partition_index,
                                   pll_partition_t * partition, int params_to_optimize,
                                   int gamma_mode, double alpha,
                                   const unsigned int * param_indices,
                                   const int * subst_matrix_symmetries);
int pllmod_treeinfo_set_active_partition(pllmod_treeinfo_t * treeinfo, int partition_index);
int pllmod_treeinfo_set_root(pllmod_treeinfo_t * treeinfo, pll_unode_t * root);
int pllmod_treeinfo_get_branch_length_all(const pllmod_treeinfo_t * treeinfo,
                                          const pll_unode_t * edge, double * lengths);
int pllmod_treeinfo_set_branch_length(pllmod_treeinfo_t * treeinfo, pll_unode_t * edge, double length);
int pllmod_treeinfo_set_branch_length_all(pllmod_treeinfo_t * treeinfo, pll_unode_t * edge,
                                          const double * lengths);
int pllmod_treeinfo_set_branch_length_partition(pllmod_treeinfo_t * treeinfo, pll_unode_t * edge,
                                                int partition_index, double length);
pllmod_treeinfo_topology_t * pllmod_treeinfo_get_topology(const pllmod_treeinfo_t * treeinfo,
                                                          pllmod_treeinfo_topology_t * topol);
int pllmod_treeinfo_set_topology(pllmod_treeinfo_t * treeinfo, const pllmod_treeinfo_topology_t * topol);
int pllmod_treeinfo_destroy_topology(pllmod_treeinfo_topology_t * topol);
int pllmod_treeinfo_destroy_partition(pllmod_treeinfo_t * treeinfo, unsigned int partition_index);
void pllmod_treeinfo_destroy(pllmod_treeinfo_t * treeinfo);
int pllmod_treeinfo_update_prob_matrices(pllmod_treeinfo_t * treeinfo, int update_all);
void pllmod_treeinfo_invalidate_all(pllmod_treeinfo_t * treeinfo);
int pllmod_treeinfo_validate_clvs(pllmod_treeinfo_t * treeinfo, pll_unode_t ** travbuffer,
                                  unsigned int travbuffer_size);
void pllmod_treeinfo_invalidate_pmatrix(pllmod_treeinfo_t * treeinfo, const pll_unode_t * edge);
void pllmod_treeinfo_invalidate_clv(pllmod_treeinfo_t * treeinfo, const pll_unode_t * edge);
double pllmod_treeinfo_compute_loglh(pllmod_treeinfo_t * treeinfo, int incremental);
double pllmod_treeinfo_compute_loglh_flex(pllmod_treeinfo_t * treeinfo, int incremental,
                                          int update_pmatrices);
double pllmod_treeinfo_compute_loglh_persite(pllmod_treeinfo_t * treeinfo, int incremental,
                                             double ** persite_lnl);
/* T:279-289, 788-790 (treeinfo.c:1558-1712) */
typedef struct
{
  unsigned int node_count;
  pll_unode_t ** nodes;
  unsigned int partition_count;
  unsigned int * partition_indices;
  pll_utree_t * tree;
  double ** probs;
} pllmod_ancestral_t;
pllmod_ancestral_t * pllmod_treeinfo_compute_ancestral(pllmod_treeinfo_t * treeinfo);
void pllmod_treeinfo_destroy_ancestral(pllmod_ancestral_t * ancestral);
int pllmod_treeinfo_scale_branches_all(pllmod_treeinfo_t * treeinfo, double scaler);
int pllmod_treeinfo_scale_branches_partition(pllmod_treeinfo_t * treeinfo,
                                             unsigned int partition_idx, double scaler);
int pllmod_treeinfo_normalize_brlen_scalers(pllmod_treeinfo_t * treeinfo);
int pllmod_treeinfo_set_tree(pllmod_treeinfo_t * treeinfo, pll_utree_t * tree);
/* src/tree/pll_tree.h:665, treeinfo.c:509: a clone of the tree carrying partition `partition_index`'s branch
 * lengths (unlinked mode; otherwise the common lengths).  The caller destroys it with pll_utree_destroy. */
pll_utree_t * pllmod_treeinfo_get_partition_tree(const pllmod_treeinfo_t * treeinfo, int partition_index);
/* Topological constraints (src/tree/pll_tree.h:724-785, treeinfo.c:1298-1556).  The legacy per-node group
 * vector (`set_constraint_clvmap`, and `constraint_check_spr` against it) is host logic and built; the
 * split-based machinery (pllmod_utree_splitset_*, pllmod_utree_resolve_multi: src/tree/utree_constraint.c,
 * outside the hot path) is not: `set_constraint_tree` fails with PLLMOD_ERROR_NOT_IMPLEMENTED, and the
 * remaining three behave as the reference does for a treeinfo without constraint splits (PLL_SUCCESS). */
int pllmod_treeinfo_set_constraint_clvmap(pllmod_treeinfo_t * treeinfo, const int * clv_index_map);
int pllmod_treeinfo_set_constraint_tree(pllmod_treeinfo_t * treeinfo, const pll_utree_t * cons_tree,
                                        int fast_and_dirty);
int pllmod_treeinfo_constraint_check_spr(pllmod_treeinfo_t * treeinfo, pll_unode_t * subtree,
                                         pll_unode_t * regraft_edge);
int pllmod_treeinfo_constraint_subtree_affected(pllmod_treeinfo_t * treeinfo, pll_unode_t * subtree);
int pllmod_treeinfo_constraint_update_splits(pllmod_treeinfo_t * treeinfo);
int pllmod_treeinfo_constraint_check_current(pllmod_treeinfo_t * treeinfo);
/* ADDITIVE (BASELINE.json north_star): traversal + P-matrix refresh + CLV update, no logL
 * evaluation and no host synchronisation; the work that treeinfo.c:973-1041 does inline. */
int pllmod_treeinfo_update_partials(pllmod_treeinfo_t * treeinfo, int incremental);

/* ---- tree rearrangement subset (T:42-64,139-189,293-330; libpll utree_moves.c:119-260) ------- */
#define PLLMOD_TREE_ERROR_SPR_INVALID_NODE 3104
#define PLLMOD_TREE_ERROR_SPR_MASK 3168
#define PLLMOD_TREE_REARRANGE_SPR 0
#define PLLMOD_TREE_REARRANGE_NNI 1
#define PLLMOD_TREE_REARRANGE_TBR 2
#define PLL_ERROR_SPR_TERMINALBRANCH 123
#define PLL_ERROR_SPR_NOCHANGE 124

/* T:139-154 */
typedef struct pll_tree_edge
{
  union
  {
    struct
    {
      pll_unode_t * parent;
      pll_unode_t * child;
    } utree;
    struct
    {
      void * parent;
      void * child;
    } rtree;
  } edge;
  double length;
} pll_tree_edge_t;

/* T:156-189 */
typedef struct
{
  int rearrange_type;
  int rooted;
  double likelihood;
  union
  {
    struct
    {
      void * prune_edge;
      void * regraft_edge;
      double prune_bl;
      double prune_left_bl;
      double prune_right_bl;
      double regraft_bl;
    } SPR;
    struct
    {
      void * edge;
      double left_left_bl;
      double left_right_bl;
      double right_left_bl;
      double right_right_bl;
      double edge_bl;
      int type;
    } NNI;
    struct
    {
      void * bisect_edge;
      pll_tree_edge_t reconn_edge;
      double bisect_left_bl;
      double bisect_right_bl;
      double reconn_parent_left_bl;
      double reconn_parent_right_bl;
      double reconn_child_left_bl;
      double reconn_child_right_bl;
    } TBR;
  };
} pll_tree_rollback_t;

/* utree_operations.c:184,229,389; pll_tree.c:159,253 (SPR rollback only) */
pll_unode_t * pllmod_utree_prune(pll_unode_t * edge);
int pllmod_utree_regraft(pll_unode_t * edge, pll_unode_t * tree);
int pllmod_utree_spr(pll_unode_t * p_edge, pll_unode_t * r_edge, pll_tree_rollback_t * rollback_info);
int pllmod_tree_rollback(pll_tree_rollback_t * rollback_info);
int pllmod_utree_nodes_at_node_dist(pll_unode_t * node, pll_unode_t ** outbuffer,
                                    unsigned int * node_count, unsigned int min_distance,
                                    unsigned int max_distance);

/* ---- SPR search round (src/algorithm/pllmod_algorithm.h:41-47,205-219; algo_search.c:1039) ---- */
typedef struct cutoff_info
{
  double lh_start;
  double lh_cutoff;
  double lh_dec_sum;
  int lh_dec_count;
} cutoff_info_t;

/* One round of subtree pruning and regrafting over every inner subnode: candidate regraft edges in
 * [radius_min, radius_max] of the pruning point are scored by partial traversals (fast) or after a
 * 3-branch Newton-Raphson optimisation (thorough); improving moves are applied, the best `ntopol_keep`
 * others are re-evaluated after full branch-length optimisation.  Returns the final logL (0 on error).
 * A constraint group vector (pllmod_treeinfo_set_constraint_clvmap) is honoured; constraint trees are not built. */
double pllmod_algo_spr_round(pllmod_treeinfo_t * treeinfo, unsigned int radius_min,
                             unsigned int radius_max, unsigned int ntopol_keep, pll_bool_t thorough,
                             int brlen_opt_method, double bl_min, double bl_max, int smoothings,
                             double epsilon, cutoff_info_t * cutoff_info, double subtree_cutoff,
                             double lh_epsilon_brlen_triplet, pll_bool_t fast_clv_updates);

/* ---- model-parameter optimisation over a treeinfo (src/algorithm/pllmod_algorithm.h:49-58,123-203;
 *      pllmod_algorithm.c:729-1870): all partitions advance in lock step, every objective evaluation is
 *      one pllmod_treeinfo_compute_loglh(treeinfo, 0) on the device.  All return -logL. ---- */
typedef int (*treeinfo_param_set_cb)(pllmod_treeinfo_t * treeinfo, unsigned int part_num,
                                     const double * param_vals, unsigned int param_count);
typedef int (*treeinfo_param_get_cb)(const pllmod_treeinfo_t * treeinfo, unsigned int part_num,
                                     double * param_vals, unsigned int param_count);
double pllmod_algo_opt_onedim_treeinfo(pllmod_treeinfo_t * treeinfo, int param_to_optimize,
                                       double min_value, double max_value, double tolerance);
double pllmod_algo_opt_onedim_treeinfo_custom(pllmod_treeinfo_t * treeinfo, int param_to_optimize,
                                              treeinfo_param_get_cb params_getter,
                                              treeinfo_param_set_cb params_setter, double min_value,
                                              double max_value, double tolerance);
double pllmod_algo_opt_subst_rates_treeinfo(pllmod_treeinfo_t * treeinfo, unsigned int params_index,
                                            double min_rate, double max_rate, double bfgs_factor,
                                            double tolerance);
double pllmod_algo_opt_frequencies_treeinfo(pllmod_treeinfo_t * treeinfo, unsigned int params_index,
                                            double min_freq, double max_freq, double bfgs_factor,
                                            double tolerance);
double pllmod_algo_opt_rates_weights_treeinfo(pllmod_treeinfo_t * treeinfo, double min_rate,
                                              double max_rate, double min_brlen, double max_brlen,
                                              double bfgs_factor, double tolerance);
double pllmod_algo_opt_alpha_pinv_treeinfo(pllmod_treeinfo_t * treeinfo, unsigned int params_index,
                                           double min_alpha, double max_alpha, double min_pinv,
                                           double max_pinv, double bfgs_factor, double tolerance);
double pllmod_algo_opt_brlen_scalers_treeinfo(pllmod_treeinfo_t * treeinfo, double min_scaler,
                                              double max_scaler, double min_brlen, double max_brlen,
                                              double lh_epsilon);
double pllmod_algo_opt_brlen_treeinfo(pllmod_treeinfo_t * treeinfo, double min_brlen,
                                      double max_brlen, double lh_epsilon, int max_iters,
                                      int opt_method, int radius);

/* ---- pll-modules optimize subset (O:60-62,71-76,88-102,158-187,194-354) --------------------- */
#define PLLMOD_OPT_DEFAULT_BRANCH_LEN 0.1
#define PLLMOD_OPT_MIN_BRANCH_LEN 1.0e-4
#define PLLMOD_OPT_MAX_BRANCH_LEN 100.
#define PLLMOD_OPT_TOL_BRANCH_LEN 1.0e-4
#define PLLMOD_OPT_BLO_NEWTON_FAST 0
#define PLLMOD_OPT_BLO_NEWTON_SAFE 1
#define PLLMOD_OPT_BLO_NEWTON_FALLBACK 2
#define PLLMOD_OPT_BLO_NEWTON_GLOBAL 3
#define PLLMOD_OPT_BLO_NEWTON_OLDFAST 10
#define PLLMOD_OPT_BLO_NEWTON_OLDSAFE 11
#define PLLMOD_OPT_ERROR_NEWTON_DERIV 2210
#define PLLMOD_OPT_ERROR_NEWTON_LIMIT 2220
#define PLLMOD_OPT_ERROR_NEWTON_UNKNOWN 2230
#define PLLMOD_OPT_ERROR_NEWTON_WORSE_LK 2240
#define PLLMOD_OPT_ERROR_NEWTON_BAD_RADIUS 2250
#define PLLMOD_OPT_BRLEN_OPTIMIZE_ALL -1

/* parameter masks, O:30-43 */
#define PLLMOD_OPT_PARAM_ALL (~0)
#define PLLMOD_OPT_PARAM_SUBST_RATES (1 << 0)
#define PLLMOD_OPT_PARAM_ALPHA (1 << 1)
#define PLLMOD_OPT_PARAM_PINV (1 << 2)
#define PLLMOD_OPT_PARAM_FREQUENCIES (1 << 3)
#define PLLMOD_OPT_PARAM_BRANCHES_SINGLE (1 << 4)
#define PLLMOD_OPT_PARAM_BRANCHES_ALL (1 << 5)
#define PLLMOD_OPT_PARAM_BRANCHES_ITERATIVE (1 << 6)
#define PLLMOD_OPT_PARAM_TOPOLOGY (1 << 7)
#define PLLMOD_OPT_PARAM_FREE_RATES (1 << 8)
#define PLLMOD_OPT_PARAM_RATE_WEIGHTS (1 << 9)
#define PLLMOD_OPT_PARAM_BRANCH_LEN_SCALER (1 << 10)
#define PLLMOD_OPT_PARAM_USER (1 << 16)
/* L-BFGS-B bound types, O:46-49 */
#define PLLMOD_OPT_LBFGSB_BOUND_NONE 0
#define PLLMOD_OPT_LBFGSB_BOUND_LOWER 1
#define PLLMOD_OPT_LBFGSB_BOUND_BOTH 2
#define PLLMOD_OPT_LBFGSB_BOUND_UPPER 3
/* defaults and limits, O:52-76 */
#define PLLMOD_OPT_DEFAULT_RATE_RATIO 1
#define PLLMOD_OPT_DEFAULT_FREQ_RATIO 1
#define PLLMOD_OPT_DEFAULT_PINV 0.01
#define PLLMOD_OPT_DEFAULT_ALPHA 0.5
#define PLLMOD_OPT_MIN_SUBST_RATE 1.0e-3
#define PLLMOD_OPT_MAX_SUBST_RATE 1000.
#define PLLMOD_OPT_MIN_FREQ 1.0e-3
#define PLLMOD_OPT_MAX_FREQ 100.
#define PLLMOD_OPT_MIN_ALPHA 0.0201
#define PLLMOD_OPT_MAX_ALPHA 100.
#define PLLMOD_OPT_MIN_PINV 0
#define PLLMOD_OPT_MAX_PINV 0.99
#define PLLMOD_OPT_LNL_UNLIKELY -1e+80
#define PLLMOD_OPT_MIN_RATE 0.02
#define PLLMOD_OPT_MAX_RATE 100.
#define PLLMOD_OPT_MIN_RATE_WEIGHT 1.0e-3
#define PLLMOD_OPT_MAX_RATE_WEIGHT 100.
#define PLLMOD_OPT_ERROR_PARAMETER 2000
#define PLLMOD_OPT_ERROR_TAXA_MISMATCH 2010
#define PLLMOD_OPT_ERROR_SEQLEN_MISMATCH 2020
#define PLLMOD_OPT_ERROR_ALIGN_UNREADABLE 2030
#define PLLMOD_OPT_ERROR_LBFGSB_UNKNOWN 2100
#define PLLMOD_OPT_ERROR_BRENT_INIT 2310

/* O:107-133: where and how the likelihood is evaluated for the Brent / L-BFGS-B drivers */
typedef struct
{
  pll_partition_t * partition;
  pll_operation_t * operations;
  double * branch_lengths;
  unsigned int * matrix_indices;
  int rooted;
  const unsigned int * params_indices;
  union
  {
    struct
    {
      unsigned int root_clv_index;
      int scaler_index;
    } rooted_t;
    struct
    {
      unsigned int parent_clv_index;
      int parent_scaler_index;
      unsigned int child_clv_index;
      int child_scaler_index;
      unsigned int edge_pmatrix_index;
    } unrooted_t;
  } where;
  char __padding__[4];
  double alpha_value;
} pll_likelihood_info_t;

/* O:135-149 */
typedef struct
{
  pll_likelihood_info_t lk_params;
  unsigned int highest_freq_state;
  unsigned int highest_weight_state;
  unsigned int params_index;
  unsigned int which_parameters;
  int * subst_params_symmetries;
  double factr;
  double pgtol;
  double * sumtable;
} pll_optimize_options_t;

/* O:158-169 */
typedef struct
{
  pll_partition_t * partition;
  pll_unode_t * tree;
  const unsigned int * params_indices;
  double * sumtable;
  double branch_length_min;
  double branch_length_max;
  double tolerance;
  int max_newton_iters;
  int opt_method;
} pll_newton_tree_params_t;

/* O:171-193 */
typedef struct
{
  pll_unode_t * tree;
  unsigned int partition_count;
  pll_partition_t ** partitions;
  unsigned int ** params_indices;
  double ** precomp_buffers;
  double ** brlen_buffers;
  double * brlen_orig;
  double * brlen_guess;
  int * converged;
  double * brlen_scalers;
  double branch_length_min;
  double branch_length_max;
  double tolerance;
  int max_newton_iters;
  int opt_method;
  int brlen_linkage;
  void * parallel_context;
  void (*parallel_reduce_cb)(void *, double *, size_t, int);
} pll_newton_tree_params_multi_t;

/* O:198-207 (opt_algorithms.c:133) */
int pllmod_opt_minimize_newton_multi(unsigned int xnum, double xmin, double * xguess, double xmax,
                                     double tolerance, unsigned int max_iters, int * converged,
                                     void * params,
                                     void (*deriv_func)(void *, double *, double *, double *));
/* O:210-218 (opt_algorithms.c:86) */
double pllmod_opt_minimize_newton(double xmin, double xguess, double xmax, double tolerance,
                                  unsigned int max_iters, void * params,
                                  void (*deriv_func)(void *, double, double *, double *));
/* O:221-229 (opt_algorithms.c:281) */
double pllmod_opt_minimize_newton_old(double x1, double xguess, double x2, double tolerance,
                                      unsigned int max_iters, void * params,
                                      void (*deriv_func)(void *, double, double *, double *));
/* O:231-241 (opt_algorithms.c:418): L-BFGS-B with forward-difference gradients */
double pllmod_opt_minimize_lbfgsb(double * x, double * xmin, double * xmax, int * bound,
                                  unsigned int n, double factr, double pgtol, void * params,
                                  double (*target_funk)(void *, double *));
/* O:244-253 (opt_algorithms.c:1404): Brent within [xmin,xmax]; returns argmin, *fx = score */
double pllmod_opt_minimize_brent(double xmin, double xguess, double xmax, double xtol, double * fx,
                                 double * f2x, void * params, double (*target_funk)(void *, double));
/* O:256-264 (opt_algorithms.c:1473): EM for mixture weights */
void pllmod_opt_minimize_em(double * w, unsigned int w_count, double * sitecat_lh,
                            unsigned int * site_w, unsigned int l, void * params,
                            double (*update_sitecatlk_funk)(void *, double *));
/* O:269-271 (pll_optimize.c:1146): derivative callback over pll_optimize_options_t */
void pllmod_opt_derivative_func(void * parameters, double proposal, double * df, double * ddf);
/* O:274-276 (pll_optimize.c:411): alpha / pinv / single branch by Brent; returns -logL */
double pllmod_opt_optimize_onedim(pll_optimize_options_t * p, double min, double max);
/* O:278-280 (pll_optimize.c:473): any parameter subset by L-BFGS-B; returns -logL */
double pllmod_opt_optimize_multidim(pll_optimize_options_t * p, double * umin, double * umax);
/* O:324-339 (opt_algorithms.c:1448): lock-step Brent over xnum variables */
int pllmod_opt_minimize_brent_multi(unsigned int xnum, int * opt_mask, double * xmin,
                                    double * xguess, double * xmax, double xtol, double * xopt,
                                    double * fx, double * f2x, void * params,
                                    double (*target_funk)(void *, double *, double *, int *),
                                    int global_range);
/* O:341-354 (opt_algorithms.c:628): lock-step L-BFGS-B over xnum partitions */
double pllmod_opt_minimize_lbfgsb_multi(unsigned int xnum, double ** x, double ** xmin,
                                        double ** xmax, int ** bound, unsigned int * n,
                                        unsigned int nmax, double factr, double pgtol, void * params,
                                        double (*target_funk)(void *, double **, double *, int *));
/* O:283-291 (pll_optimize.c:1115) */
double pllmod_opt_optimize_branch_lengths_iterative(pll_partition_t * partition, pll_unode_t * tree,
                                                    const unsigned int * params_indices,
                                                    double branch_length_min,
                                                    double branch_length_max, double tolerance,
                                                    int smoothings, int keep_update);
/* O:293-302 (pll_optimize.c:961) */
double pllmod_opt_optimize_branch_lengths_local(pll_partition_t * partition, pll_unode_t * tree,
                                                const unsigned int * params_indices,
                                                double branch_length_min, double branch_length_max,
                                                double tolerance, int smoothings, int radius,
                                                int keep_update);
/* O:304-322 (pll_optimize.c:1739) */
double pllmod_opt_optimize_branch_lengths_local_multi(
    pll_partition_t ** partitions, size_t partition_count, pll_unode_t * tree,
    unsigned int ** params_indices, double ** sumtable_buffers, double ** brlen_buffers,
    double * brlen_scalers, double branch_length_min, double branch_length_max,
    double lh_epsilon, int max_iters, int radius, int keep_update, int opt_method,
    int brlen_linkage, void * parallel_context,
    void (*parallel_reduce_cb)(void *, double *, size_t, int));
/* pll_optimize.c:1182 */
double pllmod_opt_compute_edge_loglikelihood_multi(
    pll_partition_t ** partitions, size_t partition_count, unsigned int parent_clv_index,
    int parent_scaler_index, unsigned int child_clv_index, int child_scaler_index,
    unsigned int matrix_index, unsigned int ** const params_indices, double * persite_lnl,
    void * parallel_context, void (*parallel_reduce_cb)(void *, double *, size_t, int));

/* ---- pll-modules binary checkpoint I/O (src/binary/pll_binary.h:29-181) ----------------------
 * Same on-disk format as the reference (headers, block map, partition / CLV / tree / custom blocks), so
 * checkpoints written by either library load in the other.  CLVs, scalers and P-matrices are streamed
 * between the file and HBM; there is no host copy to keep in sync.  Site-repeat blocks are accepted
 * as no-ops exactly like the reference does for partitions without PLL_ATTRIB_SITE_REPEATS. */
#define PLLMOD_BIN_BLOCK_PARTITION 0
#define PLLMOD_BIN_BLOCK_CLV 1
#define PLLMOD_BIN_BLOCK_TREE 2
#define PLLMOD_BIN_BLOCK_CUSTOM 3
#define PLLMOD_BIN_BLOCK_REPEATS 4
#define PLLMOD_BIN_ACCESS_SEQUENTIAL 0
#define PLLMOD_BIN_ACCESS_RANDOM 1
#define PLLMOD_BIN_ACCESS_SEEK -1
#define PLLMOD_BIN_INVALID_OFFSET -1
#define PLLMOD_BIN_ATTRIB_UPDATE_MAP (1 << 0)
#define PLLMOD_BIN_ATTRIB_PARTITION_DUMP_CLV (1 << 1)
#define PLLMOD_BIN_ATTRIB_PARTITION_DUMP_WGT (1 << 2)
#define PLLMOD_BIN_ATTRIB_ALIGNED (1 << 3)
#define PLLMOD_BIN_ATTRIB_PARTITION_LOAD_SKELETON (1 << 4)
#define PLLMOD_BIN_ERROR_BLOCK_MISMATCH 4001
#define PLLMOD_BIN_ERROR_BLOCK_LENGTH 4002
#define PLLMOD_BIN_ERROR_BINARY_IO 4003
#define PLLMOD_BIN_ERROR_INVALID_INDEX 4010
#define PLLMOD_BIN_ERROR_INVALID_SIZE 4011
#define PLLMOD_BIN_ERROR_LOADSTORE 4012
#define PLLMOD_BIN_ERROR_MISSING_BLOCK 4020

/* pll_binary.h:63-71 */
typedef struct
{
  unsigned int n_blocks;
  unsigned int max_blocks;
  unsigned int access_type;
  char pad[4];
  long map_offset;
} pll_binary_header_t;
/* pll_binary.h:74-78 */
typedef struct
{
  long block_id;
  long block_offset;
} pll_block_map_t;
/* pll_binary.h:87-95 */
typedef struct
{
  long block_id;
  unsigned int type;
  unsigned int attributes;
  unsigned int alignment;
  char pad[4];
  size_t block_len;
} pll_block_header_t;

/* pll_binary.h:97-181 */
FILE * pllmod_binary_create(const char * filename, pll_binary_header_t * header, unsigned int access_type,
                            unsigned int n_blocks);
FILE * pllmod_binary_open(const char * filename, pll_binary_header_t * header);
FILE * pllmod_binary_append_open(const char * filename, pll_binary_header_t * header);
int pllmod_binary_close(FILE * bin_file);
pll_block_map_t * pllmod_binary_get_map(FILE * bin_file, unsigned int * n_blocks);
int pllmod_binary_partition_dump(FILE * bin_file, int block_id, pll_partition_t * partition,
                                 unsigned int attributes);
pll_partition_t * pllmod_binary_partition_load(FILE * bin_file, int block_id, pll_partition_t * partition,
                                               unsigned int * attributes, long int offset);
int pllmod_binary_repeats_dump(FILE * bin_file, int block_id, pll_partition_t * partition,
                               unsigned int attributes);
int pllmod_binary_repeats_load(FILE * bin_file, int block_id, pll_partition_t * partition,
                               unsigned int * attributes, long int offset);
int pllmod_binary_pernoderepeats_dump(FILE * bin_file, int block_id, pll_partition_t * partition,
                                      unsigned int clv_index, unsigned int attributes);
int pllmod_binary_pernoderepeats_load(FILE * bin_file, int block_id, pll_partition_t * partition,
                                      unsigned int clv_index, unsigned int * attributes, long int offset);
int pllmod_binary_clv_dump(FILE * bin_file, int block_id, pll_partition_t * partition,
                           unsigned int clv_index, unsigned int attributes);
int pllmod_binary_clv_load(FILE * bin_file, int block_id, pll_partition_t * partition,
                           unsigned int clv_index, unsigned int * attributes, long int offset);
int pllmod_binary_utree_dump(FILE * bin_file, int block_id, pll_unode_t * tree, unsigned int tip_count,
                             unsigned int attributes);
pll_unode_t * pllmod_binary_utree_load(FILE * bin_file, int block_id, unsigned int * attributes,
                                       long int offset);
int pllmod_binary_custom_dump(FILE * bin_file, int block_id, void * data, size_t size,
                              unsigned int attributes);
void * pllmod_binary_custom_load(FILE * bin_file, int block_id, size_t * size, unsigned int * type,
                                 unsigned int * attributes, long int offset);

/* ---- additive device-side API (pllb200_*) ------------------------------------------------- */

/* Number of usable CUDA devices (0 when the driver/GPU is absent). Never fails. */
int pllb200_device_count(void);
/* Device that subsequently created partitions of THIS THREAD live on (default: current device). */
int pllb200_set_device(int device);
/* Library build info string: "pllb200 <version> sm_100a". */
const char * pllb200_version(void);

/* Blocking copies of device-resident buffers back to host memory (sizes in elements as laid out by
 * the reference: clv = sites_alloc*rate_cats*states_padded doubles, scaler = sites_alloc (or
 * sites_alloc*rate_cats with PLL_ATTRIB_RATE_SCALERS) unsigned ints, pmatrix =
 * rate_cats*states*states_padded doubles). Return PLL_SUCCESS / PLL_FAILURE. */
int pllb200_download_clv(pll_partition_t * partition, unsigned int clv_index, double * dst);
int pllb200_upload_clv(pll_partition_t * partition, unsigned int clv_index, const double * src);
int pllb200_download_scaler(pll_partition_t * partition, unsigned int scaler_index, unsigned int * dst);
int pllb200_download_pmatrix(pll_partition_t * partition, unsigned int matrix_index, double * dst);
/* All P-matrices at once, reference layout [prob_matrices][rate_cats][states][states_padded]; the upload
 * also rebuilds the kernels' interleaved copies. */
int pllb200_download_pmatrices(pll_partition_t * partition, double * dst);
int pllb200_upload_pmatrices(pll_partition_t * partition, const double * src);
int pllb200_upload_scaler(pll_partition_t * partition, unsigned int scaler_index, const unsigned int * src);
/* release the device shadow of a caller-owned sumtable buffer (pll_aligned_free(key) does the same) */
void pllb200_forget_sumtable(const void * sumtable_key);
int pllb200_download_sumtable(pll_partition_t * partition, const double * sumtable_key, double * dst);
/* Re-mirror host arrays a caller modified in place (pattern_weights / invariant / tipchars). */
int pllb200_sync_host_arrays(pll_partition_t * partition);
/* Wait for all queued device work of the partition. */
int pllb200_partition_synchronize(pll_partition_t * partition);
/* Wait only for the results queued by the *_async calls (polls their completion tags in mapped host
 * memory: no stream synchronisation) and deliver them to the addresses given at queue time. */
int pllb200_partition_wait_results(pll_partition_t * partition);
/* Run the partition's kernels on a caller-provided CUDA stream (cudaStream_t passed as void*),
 * e.g. torch.cuda.current_stream().cuda_stream.  NULL restores the partition's own stream. */
int pllb200_partition_set_stream(pll_partition_t * partition, void * cuda_stream);

/* Asynchronous log-likelihood: queues traversal-independent edge logL evaluation; the value is
 * delivered to *result (host memory owned by the caller) by pllb200_partition_synchronize or any
 * later blocking call on the partition. */
int pllb200_edge_loglikelihood_async(pll_partition_t * partition,
                                     unsigned int parent_clv_index, int parent_scaler_index,
                                     unsigned int child_clv_index, int child_scaler_index,
                                     unsigned int matrix_index,
                                     const unsigned int * freqs_indices, double * result);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).  The same value as
 * pll_compute_edge_loglikelihood (and, if persite_lnl is not NULL, the same `sites` per-site log-likelihoods),
 * computed by the same kernels, but the scratch memory — per-CTA partial sums, the reduction ticket and the
 * per-site values — lives in a static device array of the library (2.1 MB per GPU, held until exit) instead of
 * buffers from cudaMalloc: the call allocates nothing on the device.  Limits: at most 262144 sites per partition
 * (larger partitions are rejected with PLL_ERROR_PARAM_INVALID); every host-side input and the model must
 * already be mirrored on the device by an earlier ordinary call (else PLL_ERROR_PARAM_INVALID); all calls on one
 * GPU share the array, so calls whose kernels overlap on the GPU (other streams, other threads) race — serialise
 * them.  Returns PLL_SUCCESS / PLL_FAILURE, the value in *loglh_out. */
int pllb200_edge_loglikelihood_noalloc(pll_partition_t * partition, unsigned int parent_clv_index,
                                       int parent_scaler_index, unsigned int child_clv_index,
                                       int child_scaler_index, unsigned int matrix_index,
                                       const unsigned int * freqs_indices, double * loglh_out,
                                       double * persite_lnl);
/* Asynchronous twin of pll_compute_likelihood_derivatives: queues the derivative pass; *d_f / *dd_f
 * (caller-owned host memory) are filled by pllb200_partition_synchronize or any later blocking call
 * on the partition.  Lets a multi-partition Newton step queue every partition before waiting once. */
int pllb200_likelihood_derivatives_async(pll_partition_t * partition, int parent_scaler_index,
                                         int child_scaler_index, double branch_length,
                                         const unsigned int * params_indices,
                                         const double * sumtable, double * d_f, double * dd_f);
/* The whole damped Newton-Raphson iteration of pllmod_opt_minimize_newton_multi (one variable) on the
 * device: one cooperative launch and one host wait per branch instead of one per iteration.
 * Returns 1 converged, 0 iteration limit reached (pll_errno = PLLMOD_OPT_ERROR_NEWTON_LIMIT, *x_out =
 * last iterate), -1 error, -2 unsupported for this partition (asc-bias, K not a power of two, ...). */
int pllb200_newton_branch(pll_partition_t * partition, const unsigned int * params_indices,
                          const double * sumtable, double xmin, double xguess, double xmax,
                          double tolerance, unsigned int max_iters, double brlen_scaler,
                          double * x_out, int * iterations);
/* The same for ONE branch length shared by several partitions (linked or scaled branch lengths,
 * pll_optimize.c:1223-1287): one launch sums s_p * d_f and s_p^2 * dd_f over all partitions in every
 * iteration.  Null entries of `partitions` are skipped; -2 if any partition is not eligible. */
int pllb200_newton_branch_multi(pll_partition_t ** partitions, unsigned int partition_count,
                                unsigned int ** params_indices, double ** sumtables,
                                const double * brlen_scalers, double xmin, double xguess, double xmax,
                                double tolerance, unsigned int max_iters, double * x_out, int * iterations);
/* Batch mode for many small partitions on one tree: between begin and end, pll_update_partials,
 * pll_update_sumtable and pll_update_prob_matrices (one matrix) on the member partitions only collect
 * their work; pllb200_batch_flush issues one launch per kernel kind (blockIdx.y = partition).  Only these
 * calls, pllb200_newton_branch_multi and flush may be used on the members in between.  begin returns 1
 * when batch mode is on, 0 when these partitions cannot be batched (nothing changes), -1 on error. */
int pllb200_batch_begin(pll_partition_t ** partitions, unsigned int count);
int pllb200_batch_flush(void);
int pllb200_batch_active(void);
int pllb200_batch_end(void);
/* The same without a host wait, for a whole smoothing pass: queues Newton on the branch and the update of
 * its P-matrix from the length left on the device; pllb200_newton_collect waits once and returns lengths
 * and statuses (0 converged, 1 iteration limit, 2 non-finite derivative).  An unconverged branch keeps
 * its old length when `revert_on_limit` (the reference's rule for the methods without a likelihood check,
 * pll_optimize.c:1524-1549); changes below 1e-10 are dropped (:1551-1560). */
int pllb200_newton_branch_queued(pll_partition_t * partition, const unsigned int * params_indices,
                                 const double * sumtable, double xmin, double host_x0, double xmax,
                                 double tolerance, unsigned int max_iters, double brlen_scaler,
                                 unsigned int matrix_index, int use_device_x0, int revert_on_limit);
int pllb200_newton_collect(pll_partition_t * partition, const unsigned int * matrix_indices,
                           unsigned int count, double * lengths, int * status);
/* Multi-GPU (one process per GPU, sites sharded): let the fused Newton kernel sum (d_f, dd_f) over the
 * ranks itself through NVLink peer memory instead of one host callback per iteration.  Every rank calls
 * exchange_create (64-byte CUDA IPC handle out), the handles are all-gathered by the caller (MPI,
 * torch.distributed, ...) and passed to exchange_connect in rank order (<= 8 ranks of one node).  With an
 * exchange in place pllmod_opt_optimize_branch_lengths_local_multi uses the fused kernel even though a
 * parallel_reduce_cb is set (the callback still serves the log-likelihood sums). */
int pllb200_exchange_create(pll_partition_t * partition, void * ipc_handle_out);
int pllb200_exchange_connect(pll_partition_t * partition, int rank, int nranks, const void * all_handles);
int pllb200_exchange_ranks(pll_partition_t * partition);
/* The collective step of the path as a library call: a `parallel_reduce_cb` (pll_tree.h:271) whose context is a
 * partition connected with pllb200_exchange_connect.  The doubles are all-reduced (op 0 sum / 1 max / 2 min)
 * by one small kernel over NVLink peer memory, combined in rank order (bit-identical on every GPU), and come
 * back through mapped host memory: no NCCL call, no copy call.  Replaces the caller-side callback of
 * src/tree/treeinfo.c:1061-1067 and src/optimize/pll_optimize.c:1218,1270-1286. */
void pllb200_exchange_reduce_cb(void * partition, double * data, size_t count, int op);
/* Device address (as integer) of the partition's most recent reduction results
 * {logL, d_f, dd_f}: lets a caller run an NCCL all-reduce on them without a host round trip. */
uint64_t pllb200_result_device_ptr(pll_partition_t * partition);

/* Kernel tuning knobs and counters (per process). */
typedef struct
{
  unsigned long long kernel_launches;   /* kernels launched by this library so far */
  unsigned long long h2d_bytes;         /* bytes staged host->device */
  unsigned long long d2h_bytes;         /* bytes copied device->host */
  unsigned long long newton_iterations; /* derivative passes run inside the fused Newton kernel */
} pllb200_stats_t;
void pllb200_get_stats(pllb200_stats_t * out);
void pllb200_reset_stats(void);
/* Experiment knobs (never needed by a caller; DESIGN.md section 7 lists them): "trav_variant", "trav_stash", "trav_short",
 * "trav_forward", "trav_reorder", "trav_ctas_per_sm", "trav_split" / "trav_split_rounds" (sub-list length, 0 = one
 * list; chunks per resident warp below which a full traversal runs as independent sub-lists), "fuse_root",
 * "fused_newton" (0 per-iteration path, 1 TMA ring /
 * tensor-core pass, 2 register version, 3 generic kernel), "queued_newton", "spr_batch", "scratch_rows",
 * "generic_dmma" (17..64 states on the FP64 tensor cores), "fast_evals" (0 general logL / derivative kernels,
 * 1 TMA ring, 2 register tile kernels), "eval_timing".  Returns the previous value, -1 if the name is unknown. */
int pllb200_set_option(const char * name, int value);
/* Measurement aid (no reference counterpart): with option "eval_timing" = 1 the logL / derivative launchers bracket
 * their kernel with CUDA events on its stream; this returns the device time of the most recent one in microseconds
 * (-1 if none).  bench.py reports it next to the host wall-clock rate of the same calls. */
double pllb200_last_eval_kernel_us(void);

#ifdef __cplusplus
}
#endif
#endif /* PLLB200_H_ */
