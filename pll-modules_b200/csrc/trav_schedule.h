/*
 * trav_schedule.h — host-side scheduling of an operation list for the 4-state traversal kernel
 * (k_traversal_dna.cu): evaluation order, and where every child of every operation comes from.
 *
 * An operation list (pll_operation_t[], utree.c:317-366) is an expression forest.  The kernel walks it
 * chunk of sites by chunk of sites, so an inner child can be taken from
 *   - the registers that still hold the previous operation's parent            (SRC_FWD),
 *   - one of `slots` per-warp shared-memory stash slots written by its producer (SRC_STASH),
 *   - global memory                                                            (SRC_GLOBAL).
 * Ordering the forest Sethi-Ullman style (the child that needs more live values first) makes every
 * subtree whose Strahler number is <= slots + 1 run without reading a single CLV back: at BASELINE C2
 * (1000 taxa) that is all but a handful of operations.
 */
#ifndef PLLB200_TRAV_SCHEDULE_H_
#define PLLB200_TRAV_SCHEDULE_H_

#include "../../include/pllb200.h"

#include <vector>

/* 32-byte resolved operation: buffer ROW indices (the kernel multiplies by the row stride itself) */
struct __attribute__((aligned(16))) LeanOp
{
  unsigned int p;   /* parent CLV row (node index - clv_first) */
  unsigned int c1;  /* child 1: CLV row, or tip index when its source is SRC_TIP */
  unsigned int c2;
  unsigned int ps;  /* scaler rows (valid when the matching flag says so) */
  unsigned int s1;
  unsigned int s2;
  unsigned int m1f; /* matrix index of child 1 | LOPA_* << 24 */
  unsigned int m2f; /* matrix index of child 2 | LOPB_* << 24 */
};

#define LOP_MATRIX_MASK 0x00ffffffu

enum { SRC_GLOBAL = 0, SRC_TIP = 1, SRC_STASH = 2, SRC_FWD = 3 };
/* flags A (top byte of m1f) */
#define LOPA_SRC1(f) ((f) & 3u)            /* child 1 is never SRC_FWD: the host swaps the children */
#define LOPA_SLOT1(f) (((f) >> 2) & 1u)
#define LOPA_SRC2(f) (((f) >> 3) & 3u)
#define LOPA_SLOT2(f) (((f) >> 5) & 1u)
#define LOPA_PS 0x40u                      /* the parent has a scaler row */
#define LOPA_STASH_W 0x80u                 /* also write the parent (and its scaler counts) to a stash slot */
/* flags B (top byte of m2f) */
enum { SSRC_NONE = 0, SSRC_GLOBAL = 1, SSRC_SAME = 2 }; /* SAME: from where the child's CLV comes (stash / registers) */
#define LOPB_SS1(f) ((f) & 3u)
#define LOPB_SS2(f) (((f) >> 2) & 3u)
#define LOPB_WSLOT(f) (((f) >> 4) & 1u)
/* specialised body of the kernel (0 = generic).  Set when the parent has a scaler row and every child's scaler
 * counts come from where its CLV comes from (tip: none, forwarded / stash: same, global: global) */
#define LOPB_TYPE(f) (((f) >> 5) & 7u)
enum { LTYPE_GENERIC = 0, LTYPE_TIP_TIP, LTYPE_TIP_FWD, LTYPE_STASH_FWD, LTYPE_GLOBAL_FWD, LTYPE_TIP_GLOBAL, LTYPE_GLOBAL_GLOBAL,
       LTYPE_SPECIAL /* regraft scoring: LeanOp::p = LSPECIAL_*, LeanOp::ps = candidate number */ };
enum { LSPECIAL_SETUP = 1, LSPECIAL_SCORE = 2 };

struct LeanStats
{
  unsigned int ops, tip_children, fwd_children, stash_children, global_children, stash_writes, segments;
  unsigned int max_need; /* largest Sethi-Ullman label */
};

/* Re-order a valid list (every buffer written once, never read before it is written, every value
 * consumed at most once) depth-first with the child that needs more live values first; ties: heavier
 * subtree first.  Returns false (and leaves `out` empty) when the list is not such a forest: the caller
 * keeps its order.  `by_need` false = heavier subtree first only (round-1 order, the 20-state kernel). */
bool trav_order_ops(const pll_operation_t * ops, unsigned int count, unsigned int nodes, bool by_need,
                    std::vector<pll_operation_t> & out);

/* Cut a forest into INDEPENDENT lists: every maximal subtree of at most `target` operations goes whole into one
 * of the first lists (first fit, largest first: lists of at most `target` operations that read nothing another of
 * them writes), the operations above them form the last list (the "top"), which reads the subtree roots from
 * global memory.  out = the lists back to back, each in Sethi-Ullman order; bounds = their boundaries (n + 1
 * entries for n lists).  False when the list is not a forest or needs no cut (count <= target). */
bool trav_split_ops(const pll_operation_t * ops, unsigned int count, unsigned int nodes, unsigned int target,
                    std::vector<pll_operation_t> & out, std::vector<unsigned int> & bounds);

/* Resolve `ops` (in execution order) into LeanOps.  `tips` = number of pattern tips (0 when tips own
 * CLVs), `seg_len` = operations staged per shared-memory segment (stash and forwarding do not cross a
 * segment), `slots` = stash slots (0..2).  Returns false when an index does not fit the packed format. */
bool trav_lean_resolve(const pll_operation_t * ops, unsigned int count, unsigned int tips, unsigned int clv_first,
                       unsigned int seg_len, int slots, bool forward, std::vector<LeanOp> & out, LeanStats * stats);

#endif
