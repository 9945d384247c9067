"""GPU parity tests: the CUDA path (through the C ABI, via the same-source harness) against
(a) the committed golden fixtures produced by the unmodified reference (tests/golden/ref_cases.json),
(b) the compiled reference itself (oracle/_ref) on the same seeded inputs, when it travelled.

Tolerances (BASELINE.json north_star): logL 1e-8 relative, derivatives 1e-6 relative (absolute floor
1e-9 where |value| < 1e-3), scaler counts exact away from the 2^-256 threshold.
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

from pllharness import (ATTR_PATTERN_TIP, ATTR_RATE_SCALERS, BLO_NEWTON_FAST, BLO_NEWTON_OLDFAST, BLO_NEWTON_OLDSAFE,
                        BLO_NEWTON_SAFE, Harness, have_reference, load_product_package, run_concurrently)

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_cases.json")))


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def assert_deriv_close(a, b):
    a, b = np.asarray(a), np.asarray(b)
    tol = np.where(np.abs(b) < 1e-3, 1e-9, 1e-6 * np.abs(b))
    assert np.all(np.abs(a - b) <= tol), (a, b)


def fix_args(kw):
    kw = dict(kw)
    return kw


@pytest.mark.parametrize("name", sorted(GOLD.keys()))
def test_against_golden(name):
    case = GOLD[name]
    ref = case["ref"]
    kw = fix_args(case["args"])
    with Harness("b200", **kw) as h:
        ll = h.loglh()
        assert rel(ll, ref["loglh"]) < 1e-8, (ll, ref["loglh"])
        np.testing.assert_allclose(h.partition_loglh(), ref["partition_loglh"], rtol=1e-8)
        r = h.root()
        df, ddf = h.derivatives(r, ref["deriv_brlens"])
        assert_deriv_close(df, ref["df"])
        assert_deriv_close(ddf, ref["ddf"])
        sc = [int(h.scaler(i).sum()) for i in range(kw["tips"] - 2)]
        assert sc == ref["scaler_sums"]
        info = h.node_info(r)
        assert rel(float(h.clv(info["clv"]).sum()), ref["root_clv_sum"]) < 1e-9
        np.testing.assert_allclose(h.rates(), ref["rates"], rtol=1e-13)
        np.testing.assert_allclose(h.pmatrix(0).ravel()[:16], ref["pmatrix0"], rtol=1e-9, atol=1e-14)
        blo = h.optimize_branch_lengths(method=0, radius=-1, max_iters=2, lh_epsilon=0.1)
        assert rel(blo, ref["blo_fast"]) < 1e-8, (blo, ref["blo_fast"])
        np.testing.assert_allclose(h.branch_lengths(), ref["blo_brlens"], atol=1e-5, rtol=1e-5)


needs_ref = pytest.mark.skipif(not have_reference(), reason="oracle/_ref not built")


def pair(**kw):
    return Harness("ref", **kw), Harness("b200", **kw)


@needs_ref
@pytest.mark.parametrize("kw", [
    dict(tips=16, sites=10000),                                  # BASELINE config 1
    dict(tips=7, sites=33, rate_cats=2),                         # ragged: not a multiple of anything
    dict(tips=5, sites=1, rate_cats=4),                          # single site
    dict(tips=64, sites=4097, rate_cats=8),
    dict(tips=30, sites=513, attributes=0),                      # tip CLVs instead of tip patterns
    dict(tips=100, sites=2000, brlen_scale=10.0),                # deep tree: scaling active
    dict(tips=100, sites=2000, brlen_scale=10.0, attributes=ATTR_PATTERN_TIP | ATTR_RATE_SCALERS),
    dict(tips=24, sites=400, states=20),                         # protein LG+G4, generic kernels
    dict(tips=70, sites=301, states=20, brlen_scale=14.0),       # protein, deep: DMMA kernel rescale pass
    dict(tips=9, sites=77, states=20, rate_cats=1),              # protein, single category, ragged tiles
    dict(tips=6, sites=5, states=20),                            # protein Gamma4, less than one 8-site tile
    dict(tips=11, sites=9, states=20, brlen_scale=14.0),         # one full tile + one site, rescaling
    dict(tips=20, sites=130, states=20, attributes=0),           # protein with tip CLVs (all inner-inner)
    dict(tips=40, sites=200, states=20, brlen_scale=14.0, attributes=ATTR_PATTERN_TIP | ATTR_RATE_SCALERS),
    dict(tips=12, sites=300, states=7, rate_cats=3),             # odd state count (padding)
    dict(tips=16, sites=1000, pinv=0.3),
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_clvs_scalers_logl_match_reference(kw):
    ref, dev = pair(**kw)
    with ref, dev:
        lr, ld = ref.loglh(), dev.loglh()
        assert rel(ld, lr) < 1e-8, (ld, lr)
        tips = kw["tips"]
        thr = 2.0 ** -256
        for i in range(tips - 2):
            cr, cd = ref.clv(tips + i), dev.clv(tips + i)
            sr, sd = ref.scaler(i), dev.scaler(i)
            if not np.array_equal(sr, sd):
                # only sites whose unscaled maximum sits at the threshold may differ
                bad = np.nonzero(sr != sd)[0]
                per_rate = bool(kw.get("attributes", ATTR_PATTERN_TIP) & ATTR_RATE_SCALERS)
                K = kw.get("rate_cats", 4)
                for b in bad:
                    site = b // K if per_rate else b
                    m = cr[site].max()
                    assert abs(m / thr - 1) < 1e-8 or abs(m * thr - 1) < 1e-8 or abs(m / thr / thr - 1) < 1e-8, \
                        (i, b, m)
            else:
                np.testing.assert_allclose(cd, cr, rtol=1e-10, atol=0)
        for m in range(2 * tips - 3):
            np.testing.assert_allclose(dev.pmatrix(m), ref.pmatrix(m), rtol=1e-9, atol=1e-15)


AMBIG = 1 << 40  # harness seed flag: 3 % IUPAC ambiguity codes in DNA alignments


@pytest.mark.parametrize("kw", [
    dict(tips=120, sites=1000),                               # forwarding + both stash slots + global children
    dict(tips=700, sites=97),                                 # two staged segments (> 512 ops), ragged last chunk
    dict(tips=60, sites=31),                                  # less than one 32-site chunk: guarded path only
    dict(tips=100, sites=2000, brlen_scale=10.0),             # deep tree: scaling votes fire
    dict(tips=80, sites=515, seed=AMBIG | 7),                 # ambiguity codes: slow tip path
    dict(tips=40, sites=260, attributes=0),                   # tip CLVs: no tip tables, children from memory
    dict(tips=50, sites=333, pinv=0.25),
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_dna_stash_kernel_equals_the_round1_kernel_bit_for_bit(kw):
    """trav_dna4_kernel (children from registers / shared-memory stash / tip column tables; trav_variant 0) writes
    exactly the CLVs and scaler counts of trav_dna_kernel (every inner child read back, tips as 0/1 vectors;
    trav_variant 3), whatever the number of stash slots."""
    pkg = load_product_package()
    tips = kw["tips"]
    out = {}
    try:
        for variant, stash in ((3, 2), (0, 2), (0, 1), (0, 0)):
            pkg.set_option("trav_variant", variant)
            pkg.set_option("trav_stash", stash)
            with Harness("b200", **kw) as h:
                ll = h.loglh()
                out[(variant, stash)] = (ll, [h.clv(tips + i).copy() for i in range(tips - 2)],
                                         [h.scaler(i).copy() for i in range(tips - 2)])
    finally:
        pkg.set_option("trav_variant", 0)
        pkg.set_option("trav_stash", 2)
    base = out[(3, 2)]
    for key in ((0, 2), (0, 1), (0, 0)):
        got = out[key]
        assert got[0] == base[0] or rel(got[0], base[0]) < 1e-15, (key, got[0], base[0])
        for i in range(tips - 2):
            assert np.array_equal(got[2][i], base[2][i]), (key, "scaler", i)
            assert np.array_equal(got[1][i], base[1][i]), (key, "clv", i)


@pytest.mark.parametrize("kw,target", [
    (dict(tips=300, sites=5000), 16),
    (dict(tips=700, sites=2101), 64),                          # ragged last chunk, lists of several sizes
    (dict(tips=1000, sites=1000), 128),
    (dict(tips=200, sites=3000, brlen_scale=10.0), 8),         # scaling votes fire on both sides of a cut
    (dict(tips=150, sites=777, seed=AMBIG | 5), 8),            # ambiguity codes: generic body
], ids=lambda v: str(v) if isinstance(v, int) else "-".join("%s%s" % (k[:3], x) for k, x in v.items()))
def test_dna_traversal_split_into_independent_sublists_is_bit_identical(kw, target):
    """option trav_split: a full traversal as independent subtrees of at most `target` operations (blockIdx.x =
    sub-list) whose last finisher per group of chunks walks the operations above them, in one launch, writes exactly
    the CLVs, scaler counts and logL of the single-list launch."""
    pkg = load_product_package()
    lib = pkg.load()
    tips = kw["tips"]
    out = {}
    old_split, old_rounds = pkg.set_option("trav_split", 0), pkg.set_option("trav_split_rounds", 1 << 20)
    try:
        for split in (0, target):
            pkg.set_option("trav_split", split)
            lib.pllb200_debug_split_launches.restype = C.c_ulonglong
            before = lib.pllb200_debug_split_launches()
            with Harness("b200", **kw) as h:
                ll = h.loglh()
                out[split] = (ll, [h.clv(tips + i).copy() for i in range(tips - 2)],
                              [h.scaler(i).copy() for i in range(tips - 2)], lib.pllb200_debug_split_launches() - before)
    finally:
        pkg.set_option("trav_split", old_split)
        pkg.set_option("trav_split_rounds", old_rounds)
    base, got = out[0], out[target]
    assert got[3] == 1 and base[3] == 0, (got[3], base[3])   # the split really ran
    assert got[0] == base[0]
    for i in range(tips - 2):
        assert np.array_equal(got[2][i], base[2][i]), ("scaler", i)
        assert np.array_equal(got[1][i], base[1][i]), ("clv", i)


def test_dna_partial_traversals_are_bit_stable_when_a_warp_takes_several_chunks():
    """100 k sites = 3125 chunks for 2368 resident warps: warps of the persistent grid walk more than one chunk.
    A sequence of re-rootings + incremental evaluations (what an SPR round does) must give bit-identical logL on
    the stash kernel (all list lengths), on the default routing and on the round-1 kernel reading every child
    back from memory — and again on a second pass (regression: a stale asynchronous matrix copy of the previous
    chunk once raced with the first copies of the next one)."""
    pkg = load_product_package()
    tips, sites, n = 60, 100000, 120

    def run(opts):
        for k, v in (("trav_short", 3), ("trav_variant", 0), ("trav_stash", 2), ("trav_forward", 1)):
            pkg.set_option(k, opts.get(k, v))
        out = []
        with Harness("b200", tips=tips, sites=sites) as h:
            h.loglh(0)
            rng = np.random.RandomState(5)
            for nd in rng.randint(tips, 4 * tips - 6, size=n):
                h.perturb_edge(int(nd), 0.05 + 0.1 * rng.rand())
                h.set_root(int(nd))
                out.append(h.loglh(1))
            out.append(h.loglh(0))
        return out

    try:
        ref = run({"trav_variant": 3, "trav_forward": 0})
        for opts in ({"trav_short": 0}, {"trav_short": 0}, {}, {"trav_short": 0, "trav_stash": 0}, {"trav_variant": 3}):
            got = run(opts)
            bad = [i for i, (a, b) in enumerate(zip(ref, got)) if a != b]
            assert not bad, (opts, bad[:5], [(ref[i], got[i]) for i in bad[:2]])
    finally:
        for k, v in (("trav_short", 3), ("trav_variant", 0), ("trav_stash", 2), ("trav_forward", 1)):
            pkg.set_option(k, v)


@needs_ref
def test_dna_ambiguity_codes_match_reference():
    kw = dict(tips=48, sites=700, seed=AMBIG | 3)
    ref, dev = pair(**kw)
    with ref, dev:
        lr, ld = ref.loglh(), dev.loglh()
        assert rel(ld, lr) < 1e-8, (ld, lr)
        for i in range(kw["tips"] - 2):
            np.testing.assert_allclose(dev.clv(kw["tips"] + i), ref.clv(kw["tips"] + i), rtol=1e-10, atol=0)
            assert np.array_equal(dev.scaler(i), ref.scaler(i))
        r = dev.root()
        df, ddf = dev.derivatives(r, [0.01, 0.1, 1.0])
        rf, rdf = ref.derivatives(ref.root(), [0.01, 0.1, 1.0])
        assert_deriv_close(df, rf)
        assert_deriv_close(ddf, rdf)


@pytest.mark.parametrize("kw", [
    dict(tips=150, sites=203, states=20, brlen_scale=6.0),   # > 128 operations: two staged segments, scaling active
    dict(tips=40, sites=1000, states=20),
    dict(tips=33, sites=130, states=20, attributes=0),       # tip CLVs: every child is loaded or forwarded
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_protein_forwarding_kernel_equals_the_per_operation_kernel(kw):
    """20 states, Gamma4: the kernel that hands a parent's D fragments to the next operation as A fragments
    (default; tile / warp variants 22 and 26; 40 = parent tiles leave through a TMA bulk store) writes the same CLVs and scaler counts as the kernel that reads
    every inner child from memory (variant 3) — the two run their DMMA k-steps in the same order."""
    pkg = load_product_package()
    tips = kw["tips"]
    out = {}
    for variant in (3, 0, 22, 26, 40, 50):
        pkg.set_option("aa_variant", variant)
        try:
            with Harness("b200", **kw) as h:
                ll = h.loglh(0)
                out[variant] = (ll, [h.clv(tips + i).copy() for i in range(tips - 2)],
                                [h.scaler(i).copy() for i in range(tips - 2)])
        finally:
            pkg.set_option("aa_variant", 0)
    for variant in (0, 22, 26, 40, 50):
        assert rel(out[variant][0], out[3][0]) < 1e-13, (variant, out[variant][0], out[3][0])
        for i in range(tips - 2):
            assert np.array_equal(out[variant][2][i], out[3][2][i]), (variant, i)
            np.testing.assert_allclose(out[variant][1][i], out[3][1][i], rtol=1e-13, atol=0)
    assert any(int(s.sum()) > 0 for s in out[0][2]) or kw.get("brlen_scale", 1.0) == 1.0


@needs_ref
@pytest.mark.parametrize("kw", [
    dict(tips=16, sites=5000),
    dict(tips=40, sites=1500, pinv=0.25),
    dict(tips=12, sites=300, states=20),
    dict(tips=60, sites=1000, brlen_scale=10.0, attributes=ATTR_PATTERN_TIP | ATTR_RATE_SCALERS),
    dict(tips=12, sites=203, states=61, rate_cats=2),             # sumtable as a tensor-core operation (tip + inner edges)
    dict(tips=10, sites=90, states=25, attributes=0),             # the same with tip CLVs, padded 25 -> 28
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_sumtable_and_derivatives_every_edge(kw):
    """Re-root on several edges (tip edges and inner edges), compare sumtable + d/dd at 6 lengths."""
    ref, dev = pair(**kw)
    brl = [1e-4, 1e-3, 0.01, 0.1, 1.0, 10.0]
    with ref, dev:
        nsub = kw["tips"] + 3 * (kw["tips"] - 2)
        rng = np.random.RandomState(1)
        for node in rng.choice(np.arange(kw["tips"], nsub), size=6, replace=False):
            for h in (ref, dev):
                h.set_root(int(node))
            lr, ld = ref.loglh(), dev.loglh()
            assert rel(ld, lr) < 1e-8
            root = ref.root()
            assert root == dev.root()
            dfr, ddfr = ref.derivatives(root, brl)
            dfd, ddfd = dev.derivatives(root, brl)
            assert_deriv_close(dfd, dfr)
            assert_deriv_close(ddfd, ddfr)
            # the eigenbasis order/sign is solver-specific: compare basis-invariant views
            sd, sr = np.sort(dev.sumtable(), axis=-1), np.sort(ref.sumtable(), axis=-1)
            scale = np.abs(sr).max(axis=-1, keepdims=True) + 1e-300
            assert np.max(np.abs(sd - sr) / scale) < 1e-9


@needs_ref
@pytest.mark.parametrize("method", [BLO_NEWTON_FAST, BLO_NEWTON_SAFE, BLO_NEWTON_OLDFAST, BLO_NEWTON_OLDSAFE])
def test_branch_length_optimisation_matches(method):
    kw = dict(tips=24, sites=3000)
    ref, dev = pair(**kw)
    with ref, dev:
        ref.loglh(), dev.loglh()
        a = ref.optimize_branch_lengths(method=method, radius=-1, max_iters=3, lh_epsilon=0.01)
        b = dev.optimize_branch_lengths(method=method, radius=-1, max_iters=3, lh_epsilon=0.01)
        assert rel(b, a) < 1e-8, (b, a)
        # every branch within the NR tolerance (bl_min/10, pll_optimize.c:1823-1825)
        np.testing.assert_allclose(dev.branch_lengths(), ref.branch_lengths(), atol=1e-5, rtol=1e-6)
        assert rel(dev.loglh(), ref.loglh()) < 1e-8


@needs_ref
def test_single_partition_legacy_blo_and_radius():
    kw = dict(tips=20, sites=2000)
    ref, dev = pair(**kw)
    with ref, dev:
        ref.loglh(), dev.loglh()
        for radius in (0, 1, 3):
            a = ref.optimize_branch_lengths_single(radius=radius, smoothings=1, tolerance=0.1)
            b = dev.optimize_branch_lengths_single(radius=radius, smoothings=1, tolerance=0.1)
            assert rel(b, a) < 1e-8, (radius, b, a)
            np.testing.assert_allclose(dev.branch_lengths(), ref.branch_lengths(), atol=1e-5, rtol=1e-6)


@needs_ref
@pytest.mark.parametrize("kw", [
    dict(tips=30, sites=1003, states=20),                 # ragged: 1003 sites = 125 tiles of 8 + 3
    dict(tips=16, sites=5, states=20),                    # less than one tile
    dict(tips=24, sites=800, states=20, pinv=0.15),       # invariant-site term in the finishing lane
    dict(tips=60, sites=20000, states=20),                # several tiles per warp, both walking directions
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_protein_newton_kernel_matches_generic_kernel_and_reference(kw):
    """20 states, Gamma-4: newton_kernel<4> / deriv_aa4_kernel (the pass as a [sites x 80] x [80 x 3] product on the
    FP64 tensor cores, aa4_pass.cuh) against the generic one-element-per-thread kernels (fused_newton 3) and the
    reference: -logL after smoothing, every branch length, and the stand-alone derivative pass."""
    pkg = load_product_package()
    out = {}
    try:
        for mode in (1, 3):
            pkg.set_option("fused_newton", mode)
            with Harness("b200", **kw) as d:
                d.loglh()
                dd = d.derivatives(d.root(), [0.004, 0.07, 0.9])
                nll = d.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=2, lh_epsilon=0.01)
                out[mode] = (nll, d.branch_lengths().copy(), dd)
    finally:
        pkg.set_option("fused_newton", 1)
    for mode in (1,):
        assert rel(out[mode][0], out[3][0]) < 1e-12, (mode, out[mode][0], out[3][0])
        # summation order differs; NR stops at 1e-5
        np.testing.assert_allclose(out[mode][1], out[3][1], rtol=1e-7, atol=1e-10)
        np.testing.assert_allclose(out[mode][2][0], out[3][2][0], rtol=1e-10, atol=1e-9)
        np.testing.assert_allclose(out[mode][2][1], out[3][2][1], rtol=1e-10, atol=1e-9)
    with Harness("ref", **kw) as r:
        r.loglh()
        a = r.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=2, lh_epsilon=0.01)
        assert rel(out[1][0], a) < 1e-8, (out[1][0], a)
        np.testing.assert_allclose(out[1][1], r.branch_lengths(), atol=1e-5, rtol=1e-6)


@needs_ref
def test_remaining_a13_entry_points_match_reference():
    """pllmod_opt_optimize_branch_lengths_iterative, pllmod_opt_compute_edge_loglikelihood_multi and
    pllmod_opt_derivative_func (pll_optimize.c:1115,1182,1146) through the same-source harness."""
    kw = dict(tips=18, sites=[1500, 700], states=[4, 20], alpha=[0.6, 1.1])
    ref, dev = pair(**kw)
    with ref, dev:
        ref.loglh(), dev.loglh()
        for node in (ref.root(), kw["tips"] + 5):
            ref.set_root(node), dev.set_root(node)
            assert rel(dev.loglh(1), ref.loglh(1)) < 1e-8
            a, b = ref.edge_loglh_multi(node), dev.edge_loglh_multi(node)
            assert rel(b, a) < 1e-8, (node, b, a)
            assert rel(b, dev.loglh(1)) < 1e-12          # the multi-partition edge logL IS the tree logL
            for p in (0, 1):
                df, ddf = dev.derivative_func(node, [0.003, 0.05, 0.4, 2.0], p=p)
                rf, rdf = ref.derivative_func(node, [0.003, 0.05, 0.4, 2.0], p=p)
                assert_deriv_close(df, rf)
                assert_deriv_close(ddf, rdf)
                d2, dd2 = dev.derivatives(node, [0.003, 0.05, 0.4, 2.0], p=p)   # same numbers as the direct call
                assert np.array_equal(df, d2) and np.array_equal(ddf, dd2)
    kw = dict(tips=22, sites=2500)
    ref, dev = pair(**kw)
    with ref, dev:
        ref.loglh(), dev.loglh()
        a = ref.optimize_branch_lengths_iterative(smoothings=2, tolerance=1e-3)
        b = dev.optimize_branch_lengths_iterative(smoothings=2, tolerance=1e-3)
        assert rel(b, a) < 1e-8, (b, a)
        np.testing.assert_allclose(dev.branch_lengths(), ref.branch_lengths(), atol=1e-5, rtol=1e-6)


def test_partitions_driven_from_two_threads_equal_the_serial_run():
    """SURVEY.md 8b threading contract: distinct partitions (each with its own tree and treeinfo) driven from distinct
    host threads at the same time give exactly the serial results (full logL, Newton-Raphson smoothing, logL)."""
    shapes = [dict(tips=40, sites=20000), dict(tips=25, sites=4000, states=20), dict(tips=60, sites=9000, seed=5),
              dict(tips=30, sites=[3000, 2000], states=[4, 4], alpha=[0.5, 1.2])]
    hs = [Harness("b200", **kw) for kw in shapes]
    try:
        serial = run_concurrently(hs, threaded=False)
    finally:
        [h.close() for h in hs]
    for rep in range(2):
        hs = [Harness("b200", **kw) for kw in shapes]
        try:
            threaded = run_concurrently(hs, threaded=True)
        finally:
            [h.close() for h in hs]
        assert np.all(np.isfinite(threaded)) and np.all(threaded[:, 0] < 0)
        assert np.array_equal(threaded, serial), (rep, threaded, serial)


@needs_ref
def test_partial_traversals_match_full_and_reference():
    """SPR-like usage: perturb an edge, re-root there, evaluate incrementally (few ops)."""
    kw = dict(tips=50, sites=1200)
    ref, dev = pair(**kw)
    with ref, dev:
        ref.loglh(), dev.loglh()
        nsub = kw["tips"] + 3 * (kw["tips"] - 2)
        rng = np.random.RandomState(3)
        for step in range(12):
            node = int(rng.randint(kw["tips"], nsub))
            newlen = float(rng.uniform(0.01, 0.5))
            for h in (ref, dev):
                h.perturb_edge(node, newlen)
                h.set_root(node)
            c0 = dev.ops_counter()
            lr, ld = ref.loglh(1), dev.loglh(1)
            assert rel(ld, lr) < 1e-8, (step, ld, lr)
            assert dev.ops_counter() - c0 < kw["tips"] - 2   # really partial
            assert rel(dev.loglh(0), ld) < 1e-10              # incremental == full


@needs_ref
def test_multi_partition_mixed_models_linked_and_unlinked():
    for linkage in (0, 2):
        kw = dict(tips=20, sites=[700, 300, 450, 64], states=[4, 20, 4, 20], alpha=[0.3, 0.6, 1.0, 2.0],
                  linkage=linkage)
        ref, dev = pair(**kw)
        with ref, dev:
            assert rel(dev.loglh(), ref.loglh()) < 1e-8
            np.testing.assert_allclose(dev.partition_loglh(), ref.partition_loglh(), rtol=1e-8)
            a = ref.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=2, lh_epsilon=0.1)
            b = dev.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=2, lh_epsilon=0.1)
            assert rel(b, a) < 1e-8, (linkage, b, a)


@needs_ref
@pytest.mark.parametrize("linkage", [0, 1])
def test_shared_branch_length_newton_over_many_partitions(linkage):
    """Linked / scaled branch lengths: one cooperative launch sums all partitions' derivatives per
    iteration (newton_multi_kernel); same lengths as the reference and as the per-iteration path."""
    pkg = load_product_package()
    kw = dict(tips=18, sites=[900, 300, 1300, 77, 520, 33], states=[4, 20, 4, 20, 4, 4],
              alpha=[0.3, 0.6, 1.0, 2.0, 0.5, 0.8], pinv=0.15, linkage=linkage)
    ref, dev = pair(**kw)
    with ref, dev:
        assert rel(dev.loglh(), ref.loglh()) < 1e-8
        if linkage == 1:   # give the partitions different scalers first
            a = ref.algo_opt_treeinfo("brlen_scalers", 1e-2)
            b = dev.algo_opt_treeinfo("brlen_scalers", 1e-2)
            assert rel(b, a) < 1e-6
            np.testing.assert_allclose(dev.treeinfo_params()["brlen_scalers"],
                                       ref.treeinfo_params()["brlen_scalers"], rtol=1e-3)
        pkg.reset_stats()
        a = ref.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=3, lh_epsilon=0.01)
        b = dev.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=3, lh_epsilon=0.01)
        fused_launches = pkg.stats()["kernel_launches"]
        assert rel(b, a) < 1e-8, (linkage, b, a)
        np.testing.assert_allclose(dev.branch_lengths(), ref.branch_lengths(), atol=1e-5, rtol=1e-5)
    # the per-iteration path (one derivative launch per partition and iteration) gives the same answer
    _, dev2 = pair(**kw)
    pkg.set_option("fused_newton", 0)
    try:
        with dev2:
            dev2.loglh()
            if linkage == 1:
                dev2.algo_opt_treeinfo("brlen_scalers", 1e-2)
            pkg.reset_stats()
            c = dev2.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=3, lh_epsilon=0.01)
            assert rel(c, b) < 1e-9, (c, b)
            assert pkg.stats()["kernel_launches"] > 1.5 * fused_launches
    finally:
        pkg.set_option("fused_newton", 1)
    # batch mode off (one launch per partition and kernel): bit-identical lengths, several times the launches
    _, dev3 = pair(**kw)
    pkg.set_option("batch", 0)
    try:
        with dev3:
            dev3.loglh()
            if linkage == 1:
                dev3.algo_opt_treeinfo("brlen_scalers", 1e-2)
            pkg.reset_stats()
            d = dev3.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=3, lh_epsilon=0.01)
            assert d == b, (d, b)
            assert pkg.stats()["kernel_launches"] > 2 * fused_launches
    finally:
        pkg.set_option("batch", 1)


def test_protein_tip_tables_equal_the_dmma_path():
    """20 states: tip children by table lookup (default) and through the tensor pipe give the same CLVs, on a
    shallow and on a deep tree (scaling), after a branch-length change (tables rebuilt with the P-matrices)."""
    pkg = load_product_package()
    for kw in (dict(tips=30, sites=777, states=20), dict(tips=60, sites=250, states=20, brlen_scale=14.0)):
        vals = {}
        for mode in (1, 0):
            pkg.set_option("aa_tiptab", mode)
            try:
                with Harness("b200", **kw) as h:
                    a = h.loglh(0)
                    brl = h.branch_lengths()
                    h.set_branch_lengths(brl * 1.3)
                    b = h.loglh(0)
                    vals[mode] = (a, b, h.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=1,
                                                                  lh_epsilon=0.1))
            finally:
                pkg.set_option("aa_tiptab", 1)
        for x, y in zip(vals[0], vals[1]):
            assert rel(x, y) < 1e-12, (kw, vals)
        assert vals[1][0] != vals[1][1]


@needs_ref
def test_persite_root_and_alpha_change():
    kw = dict(tips=16, sites=900)
    ref, dev = pair(**kw)
    with ref, dev:
        vr, pr = ref.loglh_persite()
        vd, pd = dev.loglh_persite()
        assert rel(vd, vr) < 1e-8
        np.testing.assert_allclose(pd, pr, rtol=1e-9)
        root = ref.root()
        back = ref.node_info(root)["back"]
        node = root if not ref.node_info(root)["tip"] else back
        assert rel(dev.root_loglh(node), ref.root_loglh(node)) < 1e-8
        for alpha in (0.1, 2.5):
            ref.set_alpha(alpha), dev.set_alpha(alpha)
            assert rel(dev.loglh(), ref.loglh()) < 1e-8


def test_update_partials_symbol_and_no_cpu_fallback():
    pkg = load_product_package()
    assert pkg.device_count() >= 1
    with Harness("b200", tips=10, sites=500) as h:
        pkg.reset_stats()
        assert h.update_partials_only(0) == 1
        s = pkg.stats()
        assert s["kernel_launches"] >= 2          # P-matrix batch + traversal kernel
        assert h.loglh(1) < 0


def test_large_full_size_properties():
    """Size-independent properties at a size the CPU oracle would need minutes for:
    logL is additive over site slices and invariant under re-rooting."""
    kw = dict(tips=64, sites=200000)
    with Harness("b200", **kw) as h:
        full = h.loglh()
        parts = 0.0
        for r in range(4):
            with Harness("b200", rank=r, nranks=4, **kw) as s:
                parts += s.loglh()
        assert rel(parts, full) < 1e-10
        nsub = kw["tips"] + 3 * (kw["tips"] - 2)
        for node in (kw["tips"], kw["tips"] + 17, nsub - 1):
            h.set_root(node)
            assert rel(h.loglh(), full) < 1e-9


def test_baseline_c2_full_size_properties():
    """BASELINE.json configs[1] at its real size (1000 taxa x 1 M sites, 128 GB of CLVs): the CPU oracle cannot
    hold it, so parity goes through size-independent properties — logL is additive over site slices, one
    slice is checked against the reference itself, logL does not depend on the root, and a Newton-Raphson
    smoothing pass over all 1997 branches does not lose likelihood."""
    import torch
    if torch.cuda.get_device_properties(0).total_memory < 150e9:
        pytest.skip("needs the 180 GB of a B200")
    kw = dict(tips=1000, sites=1000000)
    with Harness("b200", **kw) as h:
        full = h.loglh()
        for node in (kw["tips"] + 5, kw["tips"] + 1501):
            h.set_root(node)
            assert rel(h.loglh(), full) < 1e-9
        nll = h.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=1, lh_epsilon=1.0)
        assert -nll > full
        assert rel(h.loglh(), -nll) < 1e-9
    parts = []
    for r in range(4):
        with Harness("b200", rank=r, nranks=4, **kw) as s:
            parts.append(s.loglh())
    assert rel(sum(parts), full) < 1e-10
    if have_reference():
        with Harness("ref", rank=7, nranks=40, **kw) as r, Harness("b200", rank=7, nranks=40, **kw) as d:
            assert r.sites == d.sites == [25000]
            assert rel(d.loglh(), r.loglh()) < 1e-8


def test_baseline_c3_full_size_properties():
    """BASELINE.json configs[2] at its real size (protein LG+G4, 500 taxa x 200 k sites, 64 GB of CLVs):
    additive over site slices, independent of the root, one slice against the reference."""
    import torch
    if torch.cuda.get_device_properties(0).total_memory < 100e9:
        pytest.skip("needs a B200-sized memory")
    kw = dict(tips=500, sites=200000, states=20)
    with Harness("b200", **kw) as h:
        full = h.loglh()
        h.set_root(kw["tips"] + 701)
        assert rel(h.loglh(), full) < 1e-9
    parts = []
    for r in range(2):
        with Harness("b200", rank=r, nranks=2, **kw) as s:
            parts.append(s.loglh())
    assert rel(sum(parts), full) < 1e-10
    if have_reference():
        with Harness("ref", rank=3, nranks=50, **kw) as r, Harness("b200", rank=3, nranks=50, **kw) as d:
            assert r.sites == d.sites == [4000]
            assert rel(d.loglh(), r.loglh()) < 1e-8


C4_SHAPE = dict(tips=200, sites=[20000] * 100, states=[4 if p % 2 == 0 else 20 for p in range(100)],
                alpha=[0.3 + 0.01 * p for p in range(100)])


@needs_ref
@pytest.mark.parametrize("shard", [0, 5])
def test_baseline_c4_real_shape_per_partition_logl_matches_reference(shard):
    """BASELINE.json configs[3] at its real shape: 100 partitions x 20 k sites, 200 taxa, DNA and protein alternating,
    per-partition Gamma shapes, batch mode on — one GPU's 1/8 site share (2500 sites of every partition: exactly what a
    GPU holds in the 8-GPU configuration; the whole config is 152 GB of CLVs) against the reference on the same share:
    every partition's logL to 1e-8, and root-edge derivatives of four partitions."""
    with Harness("ref", rank=shard, nranks=8, **C4_SHAPE) as r, Harness("b200", rank=shard, nranks=8, **C4_SHAPE) as d:
        assert r.sites == d.sites == [2500] * 100
        lr, ld = r.loglh(), d.loglh()
        np.testing.assert_allclose(d.partition_loglh(), r.partition_loglh(), rtol=1e-8, atol=0)
        assert rel(ld, lr) < 1e-8
        # derivatives of the root edge, partition by partition (a smoothing pass of the reference takes 100 s here)
        for p in (0, 1, 98, 99):
            df, ddf = d.derivatives(d.root(), [0.02, 0.3], p=p)
            rf, rdf = r.derivatives(r.root(), [0.02, 0.3], p=p)
            assert_deriv_close(df, rf)
            assert_deriv_close(ddf, rdf)


def test_baseline_c4_real_shape_shards_add_up():
    """The eight site shares of the real C4 shape, evaluated one after the other on this GPU, add up per partition
    to the same numbers whichever way the 20 k sites are cut (8 x 2500 against 4 x 5000)."""
    def total(nranks):
        acc = np.zeros(100)
        for rk in range(nranks):
            with Harness("b200", rank=rk, nranks=nranks, **C4_SHAPE) as d:
                d.loglh()
                acc += d.partition_loglh()
        return acc
    np.testing.assert_allclose(total(8), total(4), rtol=1e-10, atol=0)


@needs_ref
def test_baseline_c5_slice_matches_reference():
    """BASELINE.json configs[4] (DNA, 2000 taxa x 500 k sites): one 12.5 k-site slice of that alignment against the
    reference — full logL, partial traversals after re-rooting as the SPR search does, one edge's derivatives."""
    kw = dict(tips=2000, sites=500000)
    with Harness("ref", rank=11, nranks=40, **kw) as r, Harness("b200", rank=11, nranks=40, **kw) as d:
        assert r.sites == d.sites == [12500]
        assert rel(d.loglh(), r.loglh()) < 1e-8
        rng = np.random.RandomState(9)
        for nd in rng.randint(kw["tips"], 4 * kw["tips"] - 6, size=6):
            x = 0.05 + 0.1 * rng.rand()
            r.perturb_edge(int(nd), x), d.perturb_edge(int(nd), x)
            r.set_root(int(nd)), d.set_root(int(nd))
            assert rel(d.loglh(1), r.loglh(1)) < 1e-8
        root = d.root()
        df, ddf = d.derivatives(root, [0.01, 0.1, 1.0])
        rf, rdf = r.derivatives(r.root(), [0.01, 0.1, 1.0])
        assert_deriv_close(df, rf)
        assert_deriv_close(ddf, rdf)


@needs_ref
@pytest.mark.parametrize("kw", [
    dict(tips=12, sites=150, states=61, rate_cats=4),                  # codon-sized alphabet (row N1): generic kernels
    dict(tips=110, sites=64, states=61, rate_cats=2, brlen_scale=12.0), # deep: scaling on the generic path
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_codon_sized_alphabet_matches_reference(kw):
    """61 states (the codon alphabet size): CLVs, scalers, logL, derivatives and one smoothing pass against the
    reference's generic kernels (core_partials.c:612-764)."""
    ref, dev = pair(**kw)
    with ref, dev:
        assert rel(dev.loglh(), ref.loglh()) < 1e-8
        tips = kw["tips"]
        for i in range(tips - 2):
            assert np.array_equal(dev.scaler(i), ref.scaler(i))
            np.testing.assert_allclose(dev.clv(tips + i), ref.clv(tips + i), rtol=1e-10, atol=0)
        df, ddf = dev.derivatives(dev.root(), [0.01, 0.1, 1.0])
        rf, rdf = ref.derivatives(ref.root(), [0.01, 0.1, 1.0])
        assert_deriv_close(df, rf)
        assert_deriv_close(ddf, rdf)
        a = ref.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=1, lh_epsilon=1.0)
        b = dev.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=1, lh_epsilon=1.0)
        assert rel(b, a) < 1e-8, (b, a)


@pytest.mark.parametrize("kw", [
    dict(tips=14, sites=333, states=61),                                   # codon size, tip patterns
    dict(tips=10, sites=100, states=64, rate_cats=2),                      # full tiles
    dict(tips=12, sites=77, states=17, rate_cats=1),                       # 3 n-tiles, ragged
    dict(tips=16, sites=250, states=25, attributes=0),                     # tip CLVs: every child through the tensor pipe
    dict(tips=150, sites=130, states=32, brlen_scale=14.0),                # deep: the rescaling pass
    dict(tips=120, sites=9, states=61, brlen_scale=10.0, attributes=0),    # one tile + one site, rescaling
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_generic_dmma_kernel_equals_the_scalar_generic_kernel(kw):
    """partial_dmma_kernel (17..64 states on the FP64 tensor cores, one launch per operation) against
    partial_generic_kernel (scalar): same scaler counts, CLVs to rounding (the DMMA adds in another order)."""
    pkg = load_product_package()
    tips = kw["tips"]
    out = {}
    try:
        for mode in (1, 0):
            pkg.set_option("generic_dmma", mode)
            with Harness("b200", **kw) as h:
                ll = h.loglh(0)
                out[mode] = (ll, [h.clv(tips + i).copy() for i in range(tips - 2)],
                             [h.scaler(i).copy() for i in range(tips - 2)])
    finally:
        pkg.set_option("generic_dmma", 1)
    assert rel(out[1][0], out[0][0]) < 1e-12, (out[1][0], out[0][0])
    for i in range(tips - 2):
        assert np.array_equal(out[1][2][i], out[0][2][i]), ("scaler", i)
        np.testing.assert_allclose(out[1][1][i], out[0][1][i], rtol=1e-11, atol=0)
    if kw.get("brlen_scale", 1.0) > 1.0:
        assert any(int(s.sum()) > 0 for s in out[1][2])


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [
    dict(tips=16, sites=10000),                                   # inner-inner and tip-inner edges, full tiles
    dict(tips=9, sites=37),                                       # one tile + 5 sites
    dict(tips=5, sites=1),
    dict(tips=100, sites=2003, brlen_scale=10.0),                 # scaler counts in the site term
    dict(tips=30, sites=513, attributes=0),                       # tip CLVs
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_dna_gamma4_tile_kernels_equal_the_general_kernels(kw):
    """logl_dna4_fast_kernel / deriv_dna4_fast_kernel (32-site tiles, a lane finishes a site) against
    logl_dna_kernel<4> / deriv_dna_kernel<4>: per-site log-likelihoods bit for bit, sums to rounding, on tip edges
    and inner edges."""
    pkg = load_product_package()
    tips = kw["tips"]
    brl = [1e-4, 0.01, 0.3, 5.0]
    out = {}
    try:
        for mode in (1, 0):
            pkg.set_option("fast_evals", mode)
            with Harness("b200", **kw) as h:
                res = []
                nsub = tips + 3 * (tips - 2)
                nodes = [0, tips - 1] + list(range(tips, nsub, max(1, (nsub - tips) // 5)))
                for node in nodes:
                    h.set_root(int(node))
                    ll = h.loglh()
                    ll2, per = h.edge_loglh_multi(h.root(), persite=True)
                    df, ddf = h.derivatives(h.root(), brl)
                    res.append((ll, ll2, per.copy(), df.copy(), ddf.copy()))
                out[mode] = res
    finally:
        pkg.set_option("fast_evals", 1)
    for a, b in zip(out[1], out[0]):
        assert np.array_equal(a[2], b[2])
        assert rel(a[0], b[0]) < 1e-13 and rel(a[1], b[1]) < 1e-13
        np.testing.assert_allclose(a[3], b[3], rtol=1e-10, atol=1e-9)
        np.testing.assert_allclose(a[4], b[4], rtol=1e-10, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [
    dict(tips=14, sites=1003, states=61),                  # E = 4 x 64 = 256: two groups of 8 pieces per tile
    dict(tips=10, sites=333, states=32, rate_cats=2),      # E = 64
    dict(tips=12, sites=500, states=61, pinv=0.1),         # invariant-site term in the finishing lane
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_generic_tensor_core_newton_pass_equals_the_scalar_kernel(kw):
    """newton_kernel<5> (17..64 states: the derivative pass as a [sites x E] x [E x 3] product on the FP64 tensor
    cores, dmma_pass in aa4_pass.cuh) against newton_kernel<0> (fused_newton 2): -logL after two smoothing passes and
    every branch length."""
    pkg = load_product_package()
    out = {}
    try:
        for mode in (1, 2):
            pkg.set_option("fused_newton", mode)
            with Harness("b200", **kw) as d:
                d.loglh()
                dd = d.derivatives(d.root(), [0.003, 0.08, 1.1])   # deriv_dmma_kernel / deriv_generic_kernel
                nll = d.optimize_branch_lengths(method=BLO_NEWTON_FAST, radius=-1, max_iters=2, lh_epsilon=0.01)
                out[mode] = (nll, d.branch_lengths().copy(), dd)
    finally:
        pkg.set_option("fused_newton", 1)
    assert rel(out[1][0], out[2][0]) < 1e-12, (out[1][0], out[2][0])
    np.testing.assert_allclose(out[1][1], out[2][1], rtol=1e-7, atol=1e-10)
    np.testing.assert_allclose(out[1][2][0], out[2][2][0], rtol=1e-10, atol=1e-9)
    np.testing.assert_allclose(out[1][2][1], out[2][2][1], rtol=1e-10, atol=1e-9)


# ---- model-parameter optimisers over device-resident state ----
OPT_GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_opt.json")))


@pytest.mark.parametrize("name", sorted(OPT_GOLD["model"].keys()))
def test_model_parameter_optimisation_matches_reference(name):
    """pllmod_opt_optimize_onedim (Brent) / _multidim (L-BFGS-B): every objective call is a device
    traversal + logL.  Tolerances: final -logL 1e-7 relative (the optimisers stop at factr=1e7*eps ~ 2e-9
    relative improvements / xtol 1e-3), parameters 2e-3 relative."""
    case = OPT_GOLD["model"][name]
    with Harness("b200", **case["args"]) as h:
        start = h.loglh()
        assert rel(start, case["start_loglh"]) < 1e-8
        score = h.optimize_params(**case["opt"])
        m = h.model()
        after = h.loglh()
    assert rel(score, case["score"]) < 1e-7, (score, case["score"])
    assert rel(-after, score) < 1e-9  # the partition is left at the optimum
    # free rates + weights is a flat problem (one rate runs into its upper bound, the smallest rate is
    # barely identifiable): the score pins it, the parameters only to a few percent
    ptol = 5e-2 if "freerates" in name else 2e-3
    for k, v in case["model"].items():
        np.testing.assert_allclose(m[k], v, rtol=ptol, atol=1e-5, err_msg=k)
    if have_reference():
        with Harness("ref", **case["args"]) as r:
            rscore = r.optimize_params(**case["opt"])
        assert rel(score, rscore) < 1e-7


# ---- SPR round over device-resident CLVs (BASELINE configs[4]) ----
@pytest.mark.parametrize("batch", [1, 0], ids=["batched_scoring", "sequential_scoring"])
@pytest.mark.parametrize("name", sorted(OPT_GOLD["spr"].keys()))
def test_spr_round_ends_on_the_reference_topology(name, batch):
    """pllmod_algo_spr_round: identical final topology (every bipartition; RF distance 0), logL within the
    reference's own 1e-6 absolute assert (algo_search.c:1453), branch lengths within the NR tolerance, identical
    cutoff statistics (= the same candidate moves were visited in the same order) — with the candidates of a
    pruned subtree scored in one launch (spr_batch 1, the default; eligible: fast mode, one DNA Gamma-4 partition)
    and one after the other as the reference does (spr_batch 0), where the number of CLV operations must also
    be the reference's."""
    case = OPT_GOLD["spr"][name]
    pkg = load_product_package()
    pkg.set_option("spr_batch", batch)
    try:
        with Harness("b200", **case["args"]) as h:
            start = h.loglh()
            assert rel(start, case["start_loglh"]) < 1e-8
            pkg.reset_stats()
            lh, co = h.spr_round(**case["spr"])
            launches = pkg.stats()["kernel_launches"]
            hs, ln = h.topology_signature()
            ops = h.ops_counter()
    finally:
        pkg.set_option("spr_batch", 1)
    assert lh != 0.0
    assert [str(int(x)) for x in hs] == case["splits"], "topology differs from the reference"
    assert abs(lh - case["loglh"]) <= 1e-6 * max(1.0, abs(case["loglh"]) * 1e-3), (lh, case["loglh"])
    np.testing.assert_allclose(ln, case["brlens"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(co, case["cutoff"], rtol=1e-6)
    eligible = (not case["spr"].get("thorough")) and case["args"].get("states", 4) == 4 \
        and not isinstance(case["args"]["sites"], list)
    if batch and eligible:
        assert ops < case["ops"], "batched scoring did not run"
        test_spr_round_ends_on_the_reference_topology.launches[name] = launches
    else:
        assert ops == case["ops"]
        if eligible and name in test_spr_round_ends_on_the_reference_topology.launches:
            # one scoring launch per pruned subtree instead of ~5 launches per candidate (the final all-branch
            # optimisation of the kept topologies, the larger part on these small trees, is the same in both)
            assert test_spr_round_ends_on_the_reference_topology.launches[name] < launches


test_spr_round_ends_on_the_reference_topology.launches = {}


def test_spr_round_batched_scoring_equals_sequential_on_a_larger_tree():
    """150 taxa x 3000 sites, radius 1..5 with the cutoff rule on: the batched candidate scoring must lead the
    round through the same moves as the sequential one (same cutoff statistics, same final topology and logL)."""
    pkg = load_product_package()
    out = {}
    try:
        for batch in (1, 0):
            pkg.set_option("spr_batch", batch)
            with Harness("b200", tips=150, sites=3000, seed=23) as h:
                h.loglh(0)
                pkg.reset_stats()
                lh, co = h.spr_round(thorough=False, radius_min=1, radius_max=5, ntopol_keep=10, smoothings=16,
                                     epsilon=1.0, bl_min=1e-6)
                hs, ln = h.topology_signature()
                out[batch] = (lh, co, [int(x) for x in hs], ln, pkg.stats()["kernel_launches"])
    finally:
        pkg.set_option("spr_batch", 1)
    assert out[1][2] == out[0][2]
    assert rel(out[1][0], out[0][0]) < 1e-10, (out[1][0], out[0][0])
    np.testing.assert_allclose(out[1][1], out[0][1], rtol=1e-9)
    np.testing.assert_allclose(out[1][3], out[0][3], rtol=1e-6, atol=1e-8)
    assert out[1][4] < out[0][4], (out[1][4], out[0][4])


# ---- treeinfo-level model optimisers, all partitions in lock step ----
@pytest.mark.parametrize("name", sorted(OPT_GOLD["algo_opt"].keys()))
def test_algo_opt_treeinfo_matches_reference(name):
    """pllmod_algo_opt_{onedim,subst_rates,frequencies,alpha_pinv,rates_weights,brlen_scalers,brlen}_treeinfo.
    Tolerances: -logL 1e-6 relative (Brent xtol 1e-3 / L-BFGS-B factr 1e7 stop well before that),
    parameters 1 % (5 % for the weakly identified free-rate mixtures)."""
    case = OPT_GOLD["algo_opt"][name]
    with Harness("b200", **case["args"]) as h:
        start = h.loglh()
        assert rel(start, case["start_loglh"]) < 1e-8
        score = h.algo_opt_treeinfo(case["what"])
        after = h.loglh()
        models = [h.model(p) for p in range(h.nparts)]
        tp = h.treeinfo_params()
        brl = h.branch_lengths()
    assert rel(score, case["score"]) < 1e-6, (score, case["score"])
    assert rel(-after, score) < 1e-9
    ptol = 5e-2 if "rates_weights" in name else 1e-2
    for m, gm in zip(models, case["models"]):
        for k, v in gm.items():
            np.testing.assert_allclose(m[k], v, rtol=ptol, atol=1e-4, err_msg=k)
    for k, v in case["treeinfo"].items():
        np.testing.assert_allclose(tp[k], v, rtol=ptol, atol=1e-4, err_msg=k)
    if not case["args"].get("linkage"):
        np.testing.assert_allclose(brl, case["brlens"], rtol=ptol, atol=1e-5)


@needs_ref
@pytest.mark.parametrize("kw", [
    dict(tips=12, sites=500),                                           # DNA, tip patterns
    dict(tips=8, sites=120, states=20),                                 # protein, 20-state tip codes
    dict(tips=10, sites=257, pinv=0.2),                                 # invariant-site array in the block
    dict(tips=9, sites=100, attributes=0),                              # tips stored as CLVs
    dict(tips=150, sites=64, brlen_scale=0.01),                         # scalers active
    # per-rate scalers: the format keeps only sites_alloc words per buffer (reference quirk), so this
    # case stays where no scaling happens and checks the block sizes and everything else
    dict(tips=10, sites=90, attributes=ATTR_PATTERN_TIP | ATTR_RATE_SCALERS),
], ids=["dna", "protein", "pinv", "no_pattern_tip", "scalers", "rate_scalers"])
def test_binary_checkpoints_cross_load(tmp_path, kw):
    """pllmod_binary_partition/clv/utree/custom dump+load (src/binary/pll_binary.c): a checkpoint written
    by either library evaluates to the writer's logL in the other, from the stored CLVs, scalers and
    P-matrices alone (nothing is recomputed after loading)."""
    import os
    from pllharness import binary_load
    ref, dev = pair(**kw)
    paths = {"ref": tmp_path / "ref.bin", "b200": tmp_path / "b200.bin"}
    with ref, dev:
        want = {"ref": ref.loglh(), "b200": dev.loglh()}
        assert rel(want["b200"], want["ref"]) < 1e-8
        length = sum(ref.branch_lengths())
        ref.binary_dump(paths["ref"])
        dev.binary_dump(paths["b200"])
    assert os.path.getsize(paths["ref"]) == os.path.getsize(paths["b200"])
    for writer, path in paths.items():
        for reader in ("ref", "b200"):
            got = binary_load(reader, path)
            assert got["blocks"] == 4 and got["tips"] == kw["tips"], (writer, reader, got)
            assert rel(got["loglh"], want[writer]) < 1e-11, (writer, reader, got, want)
            assert rel(got["loglh_after_clv_block"], want[writer]) < 1e-11, (writer, reader, got)
            assert abs(got["length"] - length) < 1e-9 and abs(got["custom_sum"] - length) < 1e-9, (writer, reader)


@needs_ref
@pytest.mark.parametrize("kw", [
    dict(tips=12, sites=300),
    dict(tips=8, sites=77, states=20),
    dict(tips=9, sites=100, attributes=0),
    dict(tips=120, sites=50, brlen_scale=0.01),                     # per-site scalers cancel
    dict(tips=10, sites=[120, 40], states=[4, 20], alpha=[0.5, 1.5]),
], ids=["dna", "protein", "no_pattern_tip", "scalers", "two_partitions"])
def test_ancestral_states_match_reference(kw):
    """pllmod_treeinfo_compute_ancestral / pll_compute_node_ancestral (treeinfo.c:1611, likelihood.c:639)."""
    ref, dev = pair(**kw)
    with ref, dev:
        ref.loglh(), dev.loglh()
        want, wl = ref.ancestral()
        got, gl = dev.ancestral()
        assert wl == gl and wl > 0                                   # Node1.. labels in the same places
        np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-12)
        s = kw["states"][0] if isinstance(kw.get("states", 4), list) else kw.get("states", 4)
        n0 = (kw["sites"][0] if isinstance(kw["sites"], list) else kw["sites"]) * s
        np.testing.assert_allclose(got[:, :n0].reshape(got.shape[0], -1, s).sum(axis=2), 1.0, rtol=1e-12)
        assert rel(dev.loglh(), ref.loglh()) < 1e-8                   # the root is back where it was
