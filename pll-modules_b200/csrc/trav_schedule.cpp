/*
 * trav_schedule.cpp — see trav_schedule.h.  Pure host code (no CUDA): unit-tested on the CPU through
 * pllb200_debug_schedule (tests/test_oracle_cpu.py).
 */
#include "trav_schedule.h"

#include <algorithm>
#include <cstring>
#include <utility>

namespace
{
struct Scratch
{
  std::vector<int> producer;  /* node -> op index, -1 */
  std::vector<unsigned int> touched;
  std::vector<int> size, need, consumer;
  std::vector<std::pair<int, int>> stack;
  /* lean_resolve */
  std::vector<int> ncons, iv_end, iv_slot;
};
thread_local Scratch t_s;

/* producer[] is kept all -1 between calls: only the entries a call touched are reset */
void producer_reset(Scratch & s)
{
  for (unsigned int n : s.touched) s.producer[n] = -1;
  s.touched.clear();
}
} // namespace

/* Reads `ops` as an expression forest: producer (node -> op), consumer, subtree size and Sethi-Ullman label of
 * every op.  False (producer[] reset) when the list is not such a forest.  On success the caller resets producer[]. */
static bool forest_build(const pll_operation_t * ops, unsigned int count, unsigned int nodes, Scratch & s)
{
  if (s.producer.size() < nodes) s.producer.resize(nodes, -1);
  bool ok = true;
  for (unsigned int i = 0; i < count && ok; ++i)
  {
    const unsigned int pn = ops[i].parent_clv_index;
    if (pn >= nodes || s.producer[pn] >= 0)
      ok = false; /* written twice: the caller's order is semantic */
    else
    {
      s.producer[pn] = (int)i;
      s.touched.push_back(pn);
    }
  }
  s.size.assign(count, 1);
  s.need.assign(count, 0);
  s.consumer.assign(count, -1);
  for (unsigned int i = 0; i < count && ok; ++i)
  {
    const unsigned int ch[2] = {ops[i].child1_clv_index, ops[i].child2_clv_index};
    for (int q = 0; q < 2 && ok; ++q)
      if (ch[q] < nodes && s.producer[ch[q]] >= 0)
      {
        const int pj = s.producer[ch[q]];
        if (pj >= (int)i || s.consumer[pj] >= 0)
          ok = false; /* reads a value written later, or a shared subtree */
        else
          s.consumer[pj] = (int)i;
      }
  }
  if (!ok)
  {
    producer_reset(s);
    return false;
  }
  /* producers precede consumers: one forward pass gives sizes and Sethi-Ullman labels */
  for (unsigned int i = 0; i < count; ++i)
  {
    int pc[2] = {-1, -1};
    const unsigned int ch[2] = {ops[i].child1_clv_index, ops[i].child2_clv_index};
    for (int q = 0; q < 2; ++q)
      if (ch[q] < nodes && s.producer[ch[q]] >= 0) pc[q] = s.producer[ch[q]];
    if (pc[0] >= 0 && pc[1] >= 0)
      s.need[i] = s.need[pc[0]] == s.need[pc[1]] ? s.need[pc[0]] + 1 : std::max(s.need[pc[0]], s.need[pc[1]]);
    else if (pc[0] >= 0 || pc[1] >= 0)
      s.need[i] = s.need[pc[0] >= 0 ? pc[0] : pc[1]];
    if (s.consumer[i] >= 0) s.size[s.consumer[i]] += s.size[i];
  }
  return true;
}

/* Appends the subtree of op `root` to `out`, depth-first, the child that needs more live values first (ties and
 * !by_need: heavier subtree first).  With `cut_below` > 0 children whose subtree has at most that many ops are
 * not entered (they belong to another list). */
static void forest_emit(const pll_operation_t * ops, unsigned int nodes, bool by_need, Scratch & s, int root,
                        int cut_below, std::vector<pll_operation_t> & out)
{
  s.stack.clear();
  s.stack.push_back({root, 0});
  while (!s.stack.empty())
  {
    auto & top = s.stack.back();
    const int i = top.first;
    if (top.second == 1)
    {
      out.push_back(ops[i]);
      s.stack.pop_back();
      continue;
    }
    top.second = 1;
    int ch[2] = {-1, -1};
    const unsigned int cc[2] = {ops[i].child1_clv_index, ops[i].child2_clv_index};
    for (int q = 0; q < 2; ++q)
      if (cc[q] < nodes && s.producer[cc[q]] >= 0 && s.size[s.producer[cc[q]]] > cut_below) ch[q] = s.producer[cc[q]];
    int first = ch[0], second = ch[1];
    if (first < 0)
      std::swap(first, second);
    else if (second >= 0)
    {
      const bool second_first = by_need && s.need[second] != s.need[first] ? s.need[second] > s.need[first]
                                                                            : s.size[second] > s.size[first];
      if (second_first) std::swap(first, second);
    }
    /* `first` is visited first => pushed last */
    if (second >= 0) s.stack.push_back({second, 0});
    if (first >= 0) s.stack.push_back({first, 0});
  }
}

bool trav_order_ops(const pll_operation_t * ops, unsigned int count, unsigned int nodes, bool by_need,
                    std::vector<pll_operation_t> & out)
{
  out.clear();
  if (count < 3) return false;
  Scratch & s = t_s;
  if (!forest_build(ops, count, nodes, s)) return false;
  out.reserve(count);
  for (unsigned int root = 0; root < count; ++root)
    if (s.consumer[root] < 0) forest_emit(ops, nodes, by_need, s, (int)root, 0, out);
  producer_reset(s);
  if (out.size() != count)
  {
    out.clear();
    return false;
  }
  return true;
}

bool trav_split_ops(const pll_operation_t * ops, unsigned int count, unsigned int nodes, unsigned int target,
                    std::vector<pll_operation_t> & out, std::vector<unsigned int> & bounds)
{
  out.clear();
  bounds.clear();
  if (target < 2 || count <= target) return false;
  Scratch & s = t_s;
  if (!forest_build(ops, count, nodes, s)) return false;
  /* maximal subtrees of at most `target` ops, largest first, first fit into lists of at most `target` ops */
  std::vector<std::pair<int, int>> roots; /* (size, op) */
  unsigned int ntop = 0;
  for (unsigned int i = 0; i < count; ++i)
  {
    if (s.size[i] > (int)target)
      ++ntop;
    else if (s.consumer[i] < 0 || s.size[s.consumer[i]] > (int)target)
      roots.push_back({s.size[i], (int)i});
  }
  if (ntop == 0)
  {
    producer_reset(s);
    return false;
  }
  std::sort(roots.begin(), roots.end(), [](const std::pair<int, int> & a, const std::pair<int, int> & b) {
    return a.first != b.first ? a.first > b.first : a.second < b.second;
  });
  std::vector<unsigned int> fill;
  std::vector<std::vector<int>> bins;
  for (const auto & r : roots)
  {
    size_t b = 0;
    while (b < bins.size() && fill[b] + (unsigned int)r.first > target) ++b;
    if (b == bins.size())
    {
      bins.emplace_back();
      fill.push_back(0);
    }
    bins[b].push_back(r.second);
    fill[b] += (unsigned int)r.first;
  }
  out.reserve(count);
  bounds.push_back(0);
  for (const auto & bin : bins)
  {
    for (int r : bin) forest_emit(ops, nodes, true, s, r, 0, out);
    bounds.push_back((unsigned int)out.size());
  }
  for (unsigned int root = 0; root < count; ++root)
    if (s.consumer[root] < 0 && s.size[root] > (int)target) forest_emit(ops, nodes, true, s, (int)root, (int)target, out);
  bounds.push_back((unsigned int)out.size());
  producer_reset(s);
  if (out.size() != count)
  {
    out.clear();
    bounds.clear();
    return false;
  }
  return true;
}

bool trav_lean_resolve(const pll_operation_t * ops, unsigned int count, unsigned int tips, unsigned int clv_first,
                       unsigned int seg_len, int slots, bool forward, std::vector<LeanOp> & out, LeanStats * stats)
{
  Scratch & s = t_s;
  out.resize(count);
  if (stats) memset(stats, 0, sizeof(*stats));
  if (slots > 2) slots = 2;
  unsigned int max_node = 0;
  for (unsigned int i = 0; i < count; ++i)
  {
    const pll_operation_t & o = ops[i];
    max_node = std::max(max_node, std::max(o.parent_clv_index, std::max(o.child1_clv_index, o.child2_clv_index)));
    if (o.child1_matrix_index > LOP_MATRIX_MASK || o.child2_matrix_index > LOP_MATRIX_MASK) return false;
    if (o.parent_clv_index < clv_first) return false;
  }
  if (s.producer.size() <= max_node) s.producer.resize((size_t)max_node + 1, -1);
  s.ncons.assign(count, 0);
  s.iv_end.assign(count, -1);  /* per producing op: index of its single stash-eligible consumer */
  s.iv_slot.assign(count, -1); /* per producing op: stash slot granted */

  /* per-child bookkeeping of the pass below: producer op (or -1) of child q of op i */
  struct ChildInfo
  {
    int prod[2];
  };
  static thread_local std::vector<ChildInfo> info;
  info.resize(count);

  for (unsigned int b = 0; b < count; b += seg_len)
  {
    const unsigned int e = std::min(count, b + seg_len);
    /* pass 1: producers and consumer counts inside the segment */
    for (unsigned int i = b; i < e; ++i)
    {
      const unsigned int ch[2] = {ops[i].child1_clv_index, ops[i].child2_clv_index};
      for (int q = 0; q < 2; ++q)
      {
        int pj = -1;
        if (ch[q] >= tips && s.producer[ch[q]] >= (int)b) pj = s.producer[ch[q]];
        info[i].prod[q] = pj;
        if (pj >= 0)
        {
          ++s.ncons[pj];
          s.iv_end[pj] = (int)i;
        }
      }
      const unsigned int pn = ops[i].parent_clv_index;
      if (s.producer[pn] < 0) s.touched.push_back(pn);
      s.producer[pn] = (int)i;
    }
    producer_reset(s);

    /* pass 2: stash slots — accept as many [producer, consumer] intervals as fit `slots` at any time.
     * Intervals arrive by start; when none is free the live interval that ends LAST is the one that goes
     * without (interval scheduling on k machines, Faigle & Nawijn): values needed soon stay on chip. */
    int live[2] = {-1, -1}; /* producing op per slot */
    for (unsigned int i = b; i < e; ++i)
    {
      for (int q = 0; q < 2; ++q)
      {
        const int pj = info[i].prod[q];
        if (pj >= 0 && s.iv_slot[pj] >= 0 && live[s.iv_slot[pj]] == pj) live[s.iv_slot[pj]] = -1;
      }
      /* forwarded consumers need no slot: the value is consumed by the very next operation */
      const int end = s.iv_end[i];
      if (slots <= 0 || s.ncons[i] != 1 || end < 0) continue;
      if (forward && end == (int)i + 1) continue;
      int slot = -1;
      for (int k = 0; k < slots; ++k)
        if (live[k] < 0)
        {
          slot = k;
          break;
        }
      if (slot < 0)
      {
        int worst = 0;
        for (int k = 1; k < slots; ++k)
          if (s.iv_end[live[k]] > s.iv_end[live[worst]]) worst = k;
        if (s.iv_end[live[worst]] > end)
        {
          s.iv_slot[live[worst]] = -1; /* retroactively: that value is written to global memory only */
          slot = worst;
        }
      }
      if (slot >= 0)
      {
        live[slot] = (int)i;
        s.iv_slot[i] = slot;
      }
    }

    /* pass 3: emit */
    for (unsigned int i = b; i < e; ++i)
    {
      const pll_operation_t & o = ops[i];
      unsigned int cidx[2] = {o.child1_clv_index, o.child2_clv_index};
      unsigned int midx[2] = {o.child1_matrix_index, o.child2_matrix_index};
      int sidx[2] = {o.child1_scaler_index, o.child2_scaler_index};
      unsigned int src[2], slot[2] = {0, 0}, ssrc[2];
      for (int q = 0; q < 2; ++q)
      {
        const int pj = info[i].prod[q];
        if (cidx[q] < tips)
          src[q] = SRC_TIP;
        else if (forward && pj == (int)i - 1 && pj >= (int)b && !(q == 1 && src[0] == SRC_FWD))
          src[q] = SRC_FWD;
        else if (pj >= 0 && s.iv_slot[pj] >= 0 && s.iv_end[pj] == (int)i && s.ncons[pj] == 1)
        {
          src[q] = SRC_STASH;
          slot[q] = (unsigned int)s.iv_slot[pj];
        }
        else
          src[q] = SRC_GLOBAL;
        if (sidx[q] < 0)
          ssrc[q] = SSRC_NONE;
        else if ((src[q] == SRC_FWD || src[q] == SRC_STASH) && ops[pj].parent_scaler_index == sidx[q])
          ssrc[q] = SSRC_SAME;
        else
          ssrc[q] = SSRC_GLOBAL;
      }
      /* the same stash slot cannot feed both children (a value has one consumer), and a forwarded value
       * always arrives as child 2, a tip next to a child from global memory as child 1: the product of the
       * two child terms commutes bit for bit */
      if (src[0] == SRC_FWD || (src[0] == SRC_GLOBAL && src[1] == SRC_TIP))
      {
        std::swap(cidx[0], cidx[1]);
        std::swap(midx[0], midx[1]);
        std::swap(sidx[0], sidx[1]);
        std::swap(src[0], src[1]);
        std::swap(slot[0], slot[1]);
        std::swap(ssrc[0], ssrc[1]);
      }
      LeanOp & d = out[i];
      d.p = o.parent_clv_index - clv_first;
      for (int q = 0; q < 2; ++q)
      {
        const unsigned int row = src[q] == SRC_TIP ? cidx[q] : cidx[q] - clv_first;
        (q ? d.c2 : d.c1) = row;
        (q ? d.s2 : d.s1) = sidx[q] < 0 ? 0u : (unsigned int)sidx[q];
      }
      d.ps = o.parent_scaler_index < 0 ? 0u : (unsigned int)o.parent_scaler_index;
      unsigned int fa = src[0] | (slot[0] << 2) | (src[1] << 3) | (slot[1] << 5);
      if (o.parent_scaler_index >= 0) fa |= LOPA_PS;
      unsigned int fb = ssrc[0] | (ssrc[1] << 2);
      if (s.iv_slot[i] >= 0)
      {
        fa |= LOPA_STASH_W;
        fb |= (unsigned int)s.iv_slot[i] << 4;
      }
      /* specialised kernel body? */
      {
        bool plain = o.parent_scaler_index >= 0;
        for (int q = 0; q < 2; ++q)
        {
          const unsigned int want = src[q] == SRC_TIP ? (unsigned int)SSRC_NONE
                                    : src[q] == SRC_GLOBAL ? (unsigned int)SSRC_GLOBAL : (unsigned int)SSRC_SAME;
          if (ssrc[q] != want) plain = false;
        }
        unsigned int type = LTYPE_GENERIC;
        if (plain)
        {
          if (src[0] == SRC_TIP && src[1] == SRC_TIP) type = LTYPE_TIP_TIP;
          else if (src[0] == SRC_TIP && src[1] == SRC_FWD) type = LTYPE_TIP_FWD;
          else if (src[0] == SRC_STASH && src[1] == SRC_FWD) type = LTYPE_STASH_FWD;
          else if (src[0] == SRC_GLOBAL && src[1] == SRC_FWD) type = LTYPE_GLOBAL_FWD;
          else if (src[0] == SRC_TIP && src[1] == SRC_GLOBAL) type = LTYPE_TIP_GLOBAL;
          else if (src[0] == SRC_GLOBAL && src[1] == SRC_GLOBAL) type = LTYPE_GLOBAL_GLOBAL;
        }
        fb |= type << 5;
      }
      d.m1f = midx[0] | (fa << 24);
      d.m2f = midx[1] | (fb << 24);
      if (stats)
      {
        for (int q = 0; q < 2; ++q)
          switch (src[q])
          {
            case SRC_TIP: ++stats->tip_children; break;
            case SRC_FWD: ++stats->fwd_children; break;
            case SRC_STASH: ++stats->stash_children; break;
            default: ++stats->global_children; break;
          }
        if (s.iv_slot[i] >= 0) ++stats->stash_writes;
      }
    }
    if (stats) ++stats->segments;
  }
  if (stats) stats->ops = count;
  return true;
}

/* CPU-testable entry: order + resolve; stats_out[8] = LeanStats.  ops_out (count LeanOps, 32 B each) may be null. */
extern "C" int pllb200_debug_schedule(const pll_operation_t * ops, unsigned int count, unsigned int nodes,
                                      unsigned int tips, unsigned int clv_first, unsigned int seg_len, int slots,
                                      int reorder, unsigned int * stats_out, void * ops_out)
{
  std::vector<pll_operation_t> ordered;
  const pll_operation_t * use = ops;
  if (reorder && trav_order_ops(ops, count, nodes, reorder == 1, ordered)) use = ordered.data();
  std::vector<LeanOp> lean;
  LeanStats st;
  if (!trav_lean_resolve(use, count, tips, clv_first, seg_len, slots, true, lean, &st)) return 0;
  if (stats_out) memcpy(stats_out, &st, sizeof(st));
  if (ops_out) memcpy(ops_out, lean.data(), sizeof(LeanOp) * count);
  return 1;
}

/* CPU-testable entry of trav_split_ops: ops_out = the re-ordered pll_operation_t list, bounds_out[0..n] = list
 * boundaries (the last list is the top).  Returns the number of lists, 0 when the list is not split. */
extern "C" int pllb200_debug_split(const pll_operation_t * ops, unsigned int count, unsigned int nodes, unsigned int target,
                                   pll_operation_t * ops_out, unsigned int * bounds_out, unsigned int max_bounds)
{
  std::vector<pll_operation_t> ordered;
  std::vector<unsigned int> bounds;
  if (!trav_split_ops(ops, count, nodes, target, ordered, bounds) || bounds.size() > max_bounds) return 0;
  memcpy(ops_out, ordered.data(), sizeof(pll_operation_t) * count);
  memcpy(bounds_out, bounds.data(), sizeof(unsigned int) * bounds.size());
  return (int)bounds.size() - 1;
}
