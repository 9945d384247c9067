"""Host-side operation scheduling of the 4-state traversal kernel (pll-modules_b200/csrc/trav_schedule.cpp):
the resolved list is replayed by a software model of the kernel's operand sources (registers of the previous
operation, two stash slots, global memory) and must compute every parent from the right children.
No GPU needed: pllb200_debug_schedule is pure host code."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from pllharness import load_product_package  # noqa: E402

SRC_GLOBAL, SRC_TIP, SRC_STASH, SRC_FWD = 0, 1, 2, 3


def random_ops(tips, seed, scalers=True, shuffle=False):
    """Post-order operation list (pll_operation_t rows) of a random rooted binary tree on `tips` leaves."""
    rng = np.random.RandomState(seed)
    roots = list(range(tips))
    nxt = tips
    ops = []
    while len(roots) > 1:
        i, j = rng.choice(len(roots), 2, replace=False)
        a, b = roots[i], roots[j]
        for x in sorted((i, j), reverse=True):
            roots.pop(x)
        p = nxt
        nxt += 1
        roots.append(p)
        sa = a - tips if (scalers and a >= tips) else -1
        sb = b - tips if (scalers and b >= tips) else -1
        # parent_clv, parent_scaler, c1_clv, c1_matrix, c1_scaler, c2_clv, c2_matrix, c2_scaler
        ops.append([p, (p - tips) if scalers else -1, a, a, sa, b, b, sb])
    ops = np.array(ops, dtype=np.int32)
    if shuffle:  # any topological order is a valid list: the joining order above already is one
        pass
    return ops, nxt


def schedule(ops, nodes, tips, seg, slots, reorder=1):
    lib = load_product_package().load()
    lib.pllb200_debug_schedule.restype = C.c_int
    lib.pllb200_debug_schedule.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int,
                                           C.c_int, C.c_void_p, C.c_void_p]
    stats = np.zeros(8, dtype=np.uint32)
    out = np.zeros((len(ops), 8), dtype=np.uint32)
    ops = np.ascontiguousarray(ops, dtype=np.int32)
    ok = lib.pllb200_debug_schedule(ops.ctypes.data, len(ops), nodes, tips, tips, seg, slots, reorder,
                                    stats.ctypes.data, out.ctypes.data)
    assert ok == 1
    names = ["ops", "tip", "fwd", "stash", "global", "stash_writes", "segments", "max_need"]
    return out, dict(zip(names, stats.tolist()))


def replay(lean, tips, seg, by_parent):
    """Software model of the kernel: symbolic values.  Returns {parent node: (child a, child b, matrices, scalers)}."""
    mem = {}        # clv row -> producing parent node (global memory)
    smem = {}       # scaler row -> producing parent
    done = {}
    for b in range(0, len(lean), seg):
        stash = [None, None]
        fw = None
        for o in range(b, min(len(lean), b + seg)):
            p, c1, c2, ps, s1, s2, m1f, m2f = [int(v) for v in lean[o]]
            fa, fb = m1f >> 24, m2f >> 24
            m1, m2 = m1f & 0xFFFFFF, m2f & 0xFFFFFF
            src = [fa & 3, (fa >> 3) & 3]
            slot = [(fa >> 2) & 1, (fa >> 5) & 1]
            ssrc = [fb & 3, (fb >> 2) & 3]
            kids, scal = [], []
            assert src[0] != SRC_FWD
            for q, (c, s) in enumerate(((c1, s1), (c2, s2))):
                if src[q] == SRC_TIP:
                    kids.append(("tip", c))
                elif src[q] == SRC_FWD:
                    assert fw is not None and fw[0] == c + tips, (o, fw, c)
                    kids.append(("clv", fw[0]))
                elif src[q] == SRC_STASH:
                    assert stash[slot[q]] is not None and stash[slot[q]][0] == c + tips, (o, stash, c)
                    kids.append(("clv", stash[slot[q]][0]))
                else:
                    kids.append(("clv", mem.get(c, ("ext", c + tips))))
                if ssrc[q] == 0:
                    scal.append(None)
                elif ssrc[q] == 1:
                    scal.append(smem.get(s, ("ext", s)))
                else:
                    holder = fw if src[q] == SRC_FWD else stash[slot[q]]
                    assert holder is not None and holder[1] == s, (o, holder, s)
                    scal.append(holder[0])
            node = p + tips
            has_ps = bool(fa & 0x40)
            done[node] = (frozenset([(kids[0], m1, scal[0]), (kids[1], m2, scal[1])]), ps if has_ps else -1)
            mem[p] = node
            if has_ps:
                smem[ps] = node
            fw = (node, ps if has_ps else -2)
            if fa & 0x80:
                stash[(fb >> 4) & 1] = (node, ps if has_ps else -2)
    return done


def expected(ops, tips):
    exp = {}
    prod_s = {}
    for p, ps, a, ma, sa, b, mb, sb in ops.tolist():
        def kid(c):
            return ("tip", c) if c < tips else ("clv", c)
        exp[p] = (frozenset([(kid(a), ma, prod_s.get(sa) if sa >= 0 else None),
                             (kid(b), mb, prod_s.get(sb) if sb >= 0 else None)]), ps)
        if ps >= 0:
            prod_s[ps] = p
    return exp


@pytest.mark.parametrize("tips,seed,seg,slots", [(16, 1, 512, 2), (100, 2, 512, 2), (1000, 3, 512, 2),
                                                 (1000, 3, 64, 2), (1000, 4, 512, 1), (1000, 5, 512, 0),
                                                 (300, 6, 7, 2), (5, 7, 512, 2), (3, 8, 512, 2)])
def test_resolved_list_replays_to_the_same_expression_forest(tips, seed, seg, slots):
    ops, nodes = random_ops(tips, seed)
    lean, st = schedule(ops, nodes, tips, seg, slots)
    got = replay(lean, tips, seg, None)
    exp = expected(ops, tips)
    assert set(got) == set(exp)
    for node in exp:
        assert got[node] == exp[node], node
    assert st["tip"] + st["fwd"] + st["stash"] + st["global"] == 2 * len(ops)
    assert st["stash"] == st["stash_writes"]


def test_sethi_ullman_order_keeps_child_reads_on_chip():
    """BASELINE C2's shape (1000 taxa): with forwarding + two stash slots only subtrees of Strahler number > 3
    read a CLV back from global memory."""
    ops, nodes = random_ops(1000, 42)
    _, st2 = schedule(ops, nodes, 1000, 512, 2)
    _, st0 = schedule(ops, nodes, 1000, 512, 0)
    _, st_size = schedule(ops, nodes, 1000, 512, 2, reorder=2)
    inner_children = 2 * len(ops) - st2["tip"]
    assert st2["global"] <= 0.05 * inner_children, st2
    assert st2["global"] <= st_size["global"], (st2, st_size)
    assert st0["stash"] == 0 and st0["global"] > st2["global"]


def test_unordered_lists_keep_their_order_and_stay_correct():
    """A list that writes a buffer twice is not re-ordered; the resolver must still be right."""
    tips = 8
    ops, nodes = random_ops(tips, 11)
    # recompute the first parent again at the end from the same children (a re-write)
    ops2 = np.vstack([ops, ops[:1]])
    lean, _ = schedule(ops2, nodes, tips, 512, 2)
    assert [int(r[0]) + tips for r in lean] == [int(r[0]) for r in ops2]
    got = replay(lean, tips, 512, None)
    exp = expected(ops2, tips)
    for node in exp:
        assert got[node] == exp[node], node


def split(ops, nodes, target):
    lib = load_product_package().load()
    lib.pllb200_debug_split.restype = C.c_int
    lib.pllb200_debug_split.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_void_p, C.c_void_p, C.c_uint]
    ops = np.ascontiguousarray(ops, dtype=np.int32)
    out = np.zeros_like(ops)
    bounds = np.zeros(len(ops) + 2, dtype=np.uint32)
    n = lib.pllb200_debug_split(ops.ctypes.data, len(ops), nodes, target, out.ctypes.data, bounds.ctypes.data, len(bounds))
    return out, bounds[:n + 1].tolist() if n else []


@pytest.mark.parametrize("tips,seed,target", [(1000, 3, 128), (1000, 4, 64), (1000, 5, 256), (300, 6, 16), (2000, 7, 128),
                                              (50, 8, 2)])
def test_split_lists_are_independent_and_cover_the_forest(tips, seed, target):
    """trav_split_ops: the first lists hold whole subtrees of at most `target` operations and may run in any order
    (here: reversed) before the top list; every list is then resolved and replayed on its own, as the kernel does."""
    ops, nodes = random_ops(tips, seed)
    out, bounds = split(ops, nodes, target)
    assert bounds and bounds[0] == 0 and bounds[-1] == len(ops)
    assert sorted(out[:, 0].tolist()) == sorted(ops[:, 0].tolist())
    lists = [out[bounds[i]:bounds[i + 1]] for i in range(len(bounds) - 1)]
    assert all(0 < len(x) <= target for x in lists[:-1]), [len(x) for x in lists]
    # no phase-1 list reads what another phase-1 list writes
    for i, a in enumerate(lists[:-1]):
        reads = set(a[:, 2].tolist()) | set(a[:, 5].tolist())
        for j, b in enumerate(lists[:-1]):
            if i != j:
                assert not (reads & set(b[:, 0].tolist()))
    # replay: phase-1 lists in reversed order, then the top, through one shared model of global memory
    mem_order = list(reversed(lists[:-1])) + [lists[-1]]
    flat = np.vstack(mem_order)
    # a concatenation replayed with one segment per list = what the two launches do
    got, exp = {}, expected(ops, tips)
    mem, smem = {}, {}
    for sub in mem_order:
        lean, _ = schedule(sub, nodes, tips, 512, 2, reorder=0)
        done = replay_shared(lean, tips, mem, smem)
        got.update(done)
    assert set(got) == set(exp) and len(flat) == len(ops)
    for node in exp:
        assert got[node] == exp[node], node
    # the top list is small: the operations with more than `target` operations below them
    if tips >= 1000:
        assert len(lists[-1]) <= 6 * len(ops) // target, (len(lists[-1]), len(lists))


def replay_shared(lean, tips, mem, smem):
    """replay() of one list against a global-memory model shared with the lists that ran before."""
    done = {}
    stash = [None, None]
    fw = None
    for o in range(len(lean)):
        p, c1, c2, ps, s1, s2, m1f, m2f = [int(v) for v in lean[o]]
        fa, fb = m1f >> 24, m2f >> 24
        m1, m2 = m1f & 0xFFFFFF, m2f & 0xFFFFFF
        src = [fa & 3, (fa >> 3) & 3]
        slot = [(fa >> 2) & 1, (fa >> 5) & 1]
        ssrc = [fb & 3, (fb >> 2) & 3]
        kids, scal = [], []
        for q, (c, s) in enumerate(((c1, s1), (c2, s2))):
            if src[q] == SRC_TIP:
                kids.append(("tip", c))
            elif src[q] == SRC_FWD:
                assert fw is not None and fw[0] == c + tips
                kids.append(("clv", fw[0]))
            elif src[q] == SRC_STASH:
                assert stash[slot[q]] is not None and stash[slot[q]][0] == c + tips
                kids.append(("clv", stash[slot[q]][0]))
            else:
                kids.append(("clv", mem.get(c, ("ext", c + tips))))
            if ssrc[q] == 0:
                scal.append(None)
            elif ssrc[q] == 1:
                scal.append(smem.get(s, ("ext", s)))
            else:
                holder = fw if src[q] == SRC_FWD else stash[slot[q]]
                assert holder is not None and holder[1] == s
                scal.append(holder[0])
        node = p + tips
        has_ps = bool(fa & 0x40)
        done[node] = (frozenset([(kids[0], m1, scal[0]), (kids[1], m2, scal[1])]), ps if has_ps else -1)
        mem[p] = node
        if has_ps:
            smem[ps] = node
        fw = (node, ps if has_ps else -2)
        if fa & 0x80:
            stash[(fb >> 4) & 1] = (node, ps if has_ps else -2)
    return done


def test_split_declines_short_lists_and_non_forests():
    ops, nodes = random_ops(40, 9)
    assert split(ops, nodes, 64)[1] == []          # nothing to cut
    ops2 = np.vstack([ops, ops[:1]])               # a buffer written twice
    assert split(ops2, nodes, 8)[1] == []
