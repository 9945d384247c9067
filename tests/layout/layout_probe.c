/* Struct-layout probe (TEST INFRASTRUCTURE): prints sizeof / offsetof of every public struct of the hot path.
 * Compiled twice by tests/test_layout_cpu.py — against include/ (the product's headers) and, where
 * /root/reference exists, against the reference's own headers — the two outputs must be identical;
 * tests/golden/struct_layout.json holds the reference's output for boxes without /root/reference. */
#include <stddef.h>
#include <stdio.h>
#include "pll.h"
#include "pll_tree.h"
#include "pll_optimize.h"
#include "pllmod_algorithm.h"
#include "pll_binary.h"

#define S(t) printf("sizeof %s %zu\n", #t, sizeof(t))
#define O(t, f) printf("offsetof %s %s %zu\n", #t, #f, offsetof(t, f))

int main(void)
{
  S(pll_partition_t);
  O(pll_partition_t, tips); O(pll_partition_t, clv_buffers); O(pll_partition_t, nodes); O(pll_partition_t, states);
  O(pll_partition_t, sites); O(pll_partition_t, pattern_weight_sum); O(pll_partition_t, rate_matrices);
  O(pll_partition_t, prob_matrices); O(pll_partition_t, rate_cats); O(pll_partition_t, scale_buffers);
  O(pll_partition_t, attributes); O(pll_partition_t, alignment); O(pll_partition_t, states_padded);
  O(pll_partition_t, clv); O(pll_partition_t, pmatrix); O(pll_partition_t, rates); O(pll_partition_t, rate_weights);
  O(pll_partition_t, subst_params); O(pll_partition_t, scale_buffer); O(pll_partition_t, frequencies);
  O(pll_partition_t, prop_invar); O(pll_partition_t, invariant); O(pll_partition_t, pattern_weights);
  O(pll_partition_t, eigen_decomp_valid); O(pll_partition_t, eigenvecs); O(pll_partition_t, inv_eigenvecs);
  O(pll_partition_t, eigenvals); O(pll_partition_t, maxstates); O(pll_partition_t, tipchars);
  O(pll_partition_t, charmap); O(pll_partition_t, ttlookup); O(pll_partition_t, tipmap);
  O(pll_partition_t, asc_bias_alloc); O(pll_partition_t, asc_additional_sites); O(pll_partition_t, repeats);
  S(pll_operation_t);
  O(pll_operation_t, parent_clv_index); O(pll_operation_t, parent_scaler_index); O(pll_operation_t, child1_clv_index);
  O(pll_operation_t, child1_matrix_index); O(pll_operation_t, child1_scaler_index); O(pll_operation_t, child2_clv_index);
  O(pll_operation_t, child2_matrix_index); O(pll_operation_t, child2_scaler_index);
  S(pll_unode_t);
  O(pll_unode_t, label); O(pll_unode_t, length); O(pll_unode_t, node_index); O(pll_unode_t, clv_index);
  O(pll_unode_t, scaler_index); O(pll_unode_t, pmatrix_index); O(pll_unode_t, next); O(pll_unode_t, back);
  O(pll_unode_t, data);
  S(pll_utree_t);
  O(pll_utree_t, tip_count); O(pll_utree_t, inner_count); O(pll_utree_t, edge_count); O(pll_utree_t, binary);
  O(pll_utree_t, nodes); O(pll_utree_t, vroot);
  S(pllmod_treeinfo_t);
  O(pllmod_treeinfo_t, root); O(pllmod_treeinfo_t, tree); O(pllmod_treeinfo_t, tip_count);
  O(pllmod_treeinfo_t, partition_count); O(pllmod_treeinfo_t, partitions); O(pllmod_treeinfo_t, alphas);
  O(pllmod_treeinfo_t, gamma_mode); O(pllmod_treeinfo_t, param_indices); O(pllmod_treeinfo_t, subst_matrix_symmetries);
  O(pllmod_treeinfo_t, branch_lengths); O(pllmod_treeinfo_t, brlen_scalers); O(pllmod_treeinfo_t, partition_loglh);
  O(pllmod_treeinfo_t, params_to_optimize); O(pllmod_treeinfo_t, constraint); O(pllmod_treeinfo_t, cons_splits);
  O(pllmod_treeinfo_t, tree_splits); O(pllmod_treeinfo_t, travbuffer); O(pllmod_treeinfo_t, matrix_indices);
  O(pllmod_treeinfo_t, operations); O(pllmod_treeinfo_t, subnodes); O(pllmod_treeinfo_t, init_partition_count);
  O(pllmod_treeinfo_t, init_partition_idx); O(pllmod_treeinfo_t, init_partitions); O(pllmod_treeinfo_t, clv_valid);
  O(pllmod_treeinfo_t, pmatrix_valid); O(pllmod_treeinfo_t, brlen_linkage); O(pllmod_treeinfo_t, linked_branch_lengths);
  O(pllmod_treeinfo_t, deriv_precomp); O(pllmod_treeinfo_t, parallel_context); O(pllmod_treeinfo_t, parallel_reduce_cb);
  S(pllmod_treeinfo_topology_t); S(pllmod_treeinfo_edge_t); S(pllmod_ancestral_t);
  S(pll_tree_rollback_t);
  O(pll_tree_rollback_t, rearrange_type); O(pll_tree_rollback_t, rooted); O(pll_tree_rollback_t, likelihood);
  O(pll_tree_rollback_t, SPR.prune_edge); O(pll_tree_rollback_t, SPR.regraft_edge); O(pll_tree_rollback_t, SPR.prune_bl);
  O(pll_tree_rollback_t, SPR.prune_left_bl); O(pll_tree_rollback_t, SPR.prune_right_bl); O(pll_tree_rollback_t, SPR.regraft_bl);
  S(pll_tree_edge_t);
  S(cutoff_info_t);
  O(cutoff_info_t, lh_start); O(cutoff_info_t, lh_cutoff); O(cutoff_info_t, lh_dec_sum); O(cutoff_info_t, lh_dec_count);
  S(pll_newton_tree_params_t);
  O(pll_newton_tree_params_t, partition); O(pll_newton_tree_params_t, tree); O(pll_newton_tree_params_t, params_indices);
  O(pll_newton_tree_params_t, sumtable); O(pll_newton_tree_params_t, branch_length_min);
  O(pll_newton_tree_params_t, branch_length_max); O(pll_newton_tree_params_t, tolerance);
  O(pll_newton_tree_params_t, max_newton_iters); O(pll_newton_tree_params_t, opt_method);
  S(pll_newton_tree_params_multi_t);
  O(pll_newton_tree_params_multi_t, partitions); O(pll_newton_tree_params_multi_t, partition_count);
  O(pll_newton_tree_params_multi_t, tree); O(pll_newton_tree_params_multi_t, params_indices);
  O(pll_newton_tree_params_multi_t, precomp_buffers); O(pll_newton_tree_params_multi_t, brlen_buffers);
  O(pll_newton_tree_params_multi_t, brlen_orig); O(pll_newton_tree_params_multi_t, brlen_guess);
  O(pll_newton_tree_params_multi_t, converged); O(pll_newton_tree_params_multi_t, brlen_scalers);
  O(pll_newton_tree_params_multi_t, branch_length_min); O(pll_newton_tree_params_multi_t, branch_length_max);
  O(pll_newton_tree_params_multi_t, tolerance); O(pll_newton_tree_params_multi_t, max_newton_iters);
  O(pll_newton_tree_params_multi_t, opt_method); O(pll_newton_tree_params_multi_t, brlen_linkage);
  O(pll_newton_tree_params_multi_t, parallel_context); O(pll_newton_tree_params_multi_t, parallel_reduce_cb);
  S(pll_optimize_options_t); S(pll_likelihood_info_t);
  S(pll_binary_header_t); S(pll_block_map_t); S(pll_block_header_t);
  return 0;
}
