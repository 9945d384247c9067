"""CPU tests (no GPU needed): the plain-C oracle restatement (oracle/pll_oracle.c) is pinned against
(a) the golden fixtures the UNMODIFIED reference produced (tests/golden/ref_cases.json) and
(b) the compiled reference itself (oracle/_ref) when present; plus host-side logic of the product
library that needs no device (Gamma rates, symbol export, loud failure without a GPU)."""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from pllharness import Harness, gamma_cats, have_reference, load_product_package, synthetic_model

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_cases.json")))
ATTR_RATE_SCALERS = 1 << 9


def oracle_lib():
    so = os.path.join(ROOT, "oracle", "libpll_oracle.so")
    src = os.path.join(ROOT, "oracle", "pll_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", src, "-o", so, "-lm"])
    lib = C.CDLL(so, mode=C.RTLD_LOCAL)
    lib.orc_run_case.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_ulonglong, C.c_double,
                                 C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_uint,
                                 C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orc_gamma_rates.argtypes = [C.c_double, C.c_uint, C.c_int, C.c_void_p]
    return lib


def run_oracle(tips, sites, states=4, rate_cats=4, attributes=1 << 4, seed=42, brlen_mean=0.1, brlen_scale=1.0,
               alpha=0.7, pinv=0.0, brl=(), site_begin=0, **_):
    lib = oracle_lib()
    rates, freqs = synthetic_model(states)
    subst = np.array(rates, dtype=np.float64)
    fr = np.array(freqs, dtype=np.float64)
    fr = fr / fr.sum() if abs(fr.sum() - 1) > 1e-8 else fr       # pll_set_frequencies normalisation
    b = np.array(brl, dtype=np.float64)
    out, df, ddf = np.zeros(1), np.zeros(len(b)), np.zeros(len(b))
    sc = np.zeros(tips - 2, dtype=np.uint64)
    ok = lib.orc_run_case(tips, sites, states, rate_cats, 1 if attributes & ATTR_RATE_SCALERS else 0, seed,
                          brlen_mean, brlen_scale, alpha, pinv, subst.ctypes.data, fr.ctypes.data, site_begin,
                          b.ctypes.data, len(b), out.ctypes.data, df.ctypes.data, ddf.ctypes.data, sc.ctypes.data)
    assert ok
    return out[0], df, ddf, sc


# the C restatement covers the core recurrences; ascertainment-bias cases are pinned by the reference's
# goldens and the live reference only (tests/test_gpu_parity.py)
SINGLE = [n for n in sorted(GOLD) if not isinstance(GOLD[n]["args"]["sites"], list) and "asc" not in n]


@pytest.mark.parametrize("name", SINGLE)
def test_oracle_matches_reference_golden(name):
    case = GOLD[name]
    ref = case["ref"]
    ll, df, ddf, sc = run_oracle(brl=ref["deriv_brlens"], **case["args"])
    assert abs(ll - ref["loglh"]) <= 1e-10 * abs(ref["loglh"]), (ll, ref["loglh"])
    np.testing.assert_allclose(df, ref["df"], rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(ddf, ref["ddf"], rtol=1e-7, atol=1e-9)
    assert [int(x) for x in sc] == ref["scaler_sums"]


def test_golden_deep_cases_exercise_scaling():
    assert sum(GOLD["dna200_deep"]["ref"]["scaler_sums"]) > 0
    assert sum(GOLD["dna200_deep_ratescalers"]["ref"]["scaler_sums"]) > 0


@pytest.mark.skipif(not have_reference(), reason="oracle/_ref not built")
@pytest.mark.parametrize("kw", [
    dict(tips=9, sites=257, states=4, rate_cats=4, seed=5),
    dict(tips=33, sites=100, states=4, rate_cats=2, seed=9, brlen_scale=12.0),
    dict(tips=33, sites=100, states=4, rate_cats=4, seed=9, brlen_scale=12.0, attributes=(1 << 4) | (1 << 9)),
    dict(tips=8, sites=64, states=20, rate_cats=4, seed=3),
    dict(tips=8, sites=50, states=3, rate_cats=1, seed=11),
    dict(tips=14, sites=120, states=4, rate_cats=4, seed=2, pinv=0.4),
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_oracle_matches_live_reference(kw):
    brl = [1e-3, 0.05, 0.7]
    with Harness("ref", **kw) as h:
        ll = h.loglh()
        df, ddf = h.derivatives(h.root(), brl)
        sc = [int(h.scaler(i).sum()) for i in range(kw["tips"] - 2)]
    oll, odf, oddf, osc = run_oracle(brl=brl, **kw)
    assert abs(oll - ll) <= 1e-10 * abs(ll)
    np.testing.assert_allclose(odf, df, rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(oddf, ddf, rtol=1e-7, atol=1e-9)
    assert [int(x) for x in osc] == sc


def test_oracle_site_slices_add_up():
    kw = dict(tips=12, sites=400, states=4, rate_cats=4, seed=8)
    full = run_oracle(**kw)[0]
    parts = 0.0
    for r in range(4):
        parts += run_oracle(**dict(kw, sites=100, site_begin=100 * r))[0]
    assert abs(parts - full) <= 1e-12 * abs(full)


# ---- host-side logic of the product that needs no device ------------------------------------------

def test_product_gamma_rates_match_oracle_and_golden():
    lib = oracle_lib()
    for alpha in (0.02, 0.1, 0.7, 1.0, 3.3, 20.0, 99.0):
        for cats in (1, 2, 4, 8, 16):
            for mode in (0, 1):
                o = np.zeros(cats)
                assert lib.orc_gamma_rates(alpha, cats, mode, o.ctypes.data)
                p = gamma_cats("b200", alpha, cats, mode)
                np.testing.assert_allclose(p, o, rtol=1e-11, atol=0)
                if have_reference():
                    np.testing.assert_allclose(p, gamma_cats("ref", alpha, cats, mode), rtol=1e-11, atol=0)
    np.testing.assert_allclose(gamma_cats("b200", 0.7, 4), GOLD["dna16"]["ref"]["rates"], rtol=1e-13)


def test_library_exports_every_declared_symbol():
    pkg = load_product_package()
    lib = pkg.load()
    declared = pkg.declared_symbols()
    assert len(declared) > 60
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    for tls in ("pll_errno", "pll_errmsg", "pll_map_nt", "pll_map_aa", "pll_map_bin"):
        out = subprocess.check_output(["nm", "-D", pkg.LIB_PATH], text=True)
        assert (" " + tls + "\n") in out, tls
    assert pkg.version().startswith("pllb200")


def test_state_maps_match_reference_tables():
    pkg = load_product_package()
    lib = pkg.load()
    nt = (C.c_ulonglong * 256).in_dll(lib, "pll_map_nt")
    aa = (C.c_ulonglong * 256).in_dll(lib, "pll_map_aa")
    assert nt[ord("A")] == 1 and nt[ord("c")] == 2 and nt[ord("G")] == 4 and nt[ord("T")] == 8
    assert nt[ord("-")] == 15 and nt[ord("N")] == 15 and nt[ord("R")] == 5 and nt[ord("Y")] == 10
    assert aa[ord("A")] == 1 and aa[ord("V")] == 1 << 19 and aa[ord("B")] == 0xC and aa[ord("-")] == 0xFFFFF
    if have_reference():
        ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libpll_ref.so"), mode=C.RTLD_LOCAL)
        for name, mine in (("pll_map_nt", nt), ("pll_map_aa", aa),
                           ("pll_map_bin", (C.c_ulonglong * 256).in_dll(lib, "pll_map_bin"))):
            theirs = (C.c_ulonglong * 256).in_dll(ref, name)
            assert list(mine) == list(theirs), name


def test_no_gpu_means_loud_failure_not_cpu_fallback():
    pkg = load_product_package()
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    lib = pkg.load()
    lib.pll_partition_create.restype = C.c_void_p
    p = lib.pll_partition_create(4, 2, 4, 10, 1, 5, 4, 2, 1 << 4)
    assert not p
    assert C.c_int.in_dll(lib, "pll_errno").value == 9001
    with pytest.raises(RuntimeError):
        Harness("b200", tips=5, sites=10)


def test_newton_minimisers_host_only():
    """pllmod_opt_minimize_newton / _old on an analytic function: f(x) = (x-2)^2 -> f' = 2(x-2), f'' = 2."""
    pkg = load_product_package()
    lib = pkg.load()
    CB = C.CFUNCTYPE(None, C.c_void_p, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double))

    def deriv(_, x, df, ddf):
        df[0] = 2 * (x - 2.0)
        ddf[0] = 2.0
    cb = CB(deriv)
    for fn in (lib.pllmod_opt_minimize_newton, lib.pllmod_opt_minimize_newton_old):
        fn.restype = C.c_double
        fn.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, C.c_uint, C.c_void_p, CB]
        x = fn(1e-4, 0.1, 100.0, 1e-6, 30, None, cb)
        assert abs(x - 2.0) < 1e-4, x
    # bounds are respected: minimum outside [xmin, xmax]
    x = lib.pllmod_opt_minimize_newton(3.0, 5.0, 100.0, 1e-6, 30, None, cb)
    assert abs(x - 3.0) < 1e-9


def test_multi_rank_site_sharding_with_gloo():
    """N>1 host logic: every rank owns a contiguous site slice, the only exchange is the all-reduce of
    the per-partition logL through parallel_reduce_cb (here: gloo, reference library as the engine since
    the CPU box has no device).  2 ranks must reproduce the single-process value."""
    if not have_reference():
        pytest.skip("oracle/_ref not built")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    script = os.path.join(ROOT, "tests", "gloo_worker.py")
    out = subprocess.check_output([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                   "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port", "29533",
                                   script], env=env, text=True, stderr=subprocess.STDOUT, timeout=300)
    line = [l for l in out.splitlines() if l.startswith("RESULT")][0]
    sharded, single = [float(x) for x in line.split()[1:3]]
    assert abs(sharded - single) <= 1e-10 * abs(single)


# ---- scalar optimisers: pure host code, so parity is checked on the CPU box ----
OPT_GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_opt.json")))


def test_brent_is_bit_identical_to_reference_golden():
    """pllmod_opt_minimize_brent (opt_algorithms.c:1404): same abscissae, same number of objective calls."""
    import pllharness as PH
    for case in OPT_GOLD["brent"]:
        r = PH.test_brent("b200", *case["args"])
        assert r["x"] == case["x"] and r["fx"] == case["fx"] and r["evals"] == case["evals"], (case, r)
        if have_reference():
            q = PH.test_brent("ref", *case["args"])
            assert q["x"] == r["x"] and q["f2x"] == r["f2x"]


def test_brent_multi_lockstep_mask_and_ranges():
    """pllmod_opt_minimize_brent_multi (opt_algorithms.c:1448): masks, per-variable ranges, global range."""
    import pllharness as PH
    a = PH.test_brent_multi("b200", [0, 1, 2, 3], [1, 1, 0, 1], [0.02], [0.5, 1.0, 2.0, 0.1], [100.0], 1e-3, 1)
    assert abs(a["x"][0] - 2.0) < 1e-3 and abs(a["x"][1] - 0.29) < 1e-3 and abs(a["x"][3] - 0.7) < 1e-3
    assert a["x"][2] == 0.0  # masked out: untouched slot, as in the reference
    b = PH.test_brent_multi("b200", [0, 2], None, [0.02, 0.1], [0.5, 1.0], [1.5, 0.6], 1e-3, 0)
    assert abs(b["x"][0] - 1.5) < 2e-3 and abs(b["x"][1] - 0.6) < 2e-3  # minima outside -> upper bounds
    if have_reference():
        ra = PH.test_brent_multi("ref", [0, 1, 2, 3], [1, 1, 0, 1], [0.02], [0.5, 1.0, 2.0, 0.1], [100.0], 1e-3, 1)
        rb = PH.test_brent_multi("ref", [0, 2], None, [0.02, 0.1], [0.5, 1.0], [1.5, 0.6], 1e-3, 0)
        assert np.array_equal(ra["x"], a["x"]) and ra["evals"] == a["evals"]
        assert np.array_equal(rb["x"], b["x"]) and rb["evals"] == b["evals"]


def test_lbfgsb_matches_reference_golden():
    """pllmod_opt_minimize_lbfgsb (opt_algorithms.c:418): our from-the-paper L-BFGS-B reproduces the
    reference's iterates on box-constrained problems (same number of objective calls; x to 1e-8), and its
    optimum within the solver's own stopping tolerance on the unconstrained one."""
    import pllharness as PH
    for case in OPT_GOLD["lbfgsb"]:
        c = case["args"]
        r = PH.test_lbfgsb("b200", c["fn"], c["x0"], c["lo"], c["hi"], c["bound"])
        assert abs(r["score"] - case["score"]) <= 1e-7 * abs(case["score"]), (case, r)
        if all(b == 2 for b in c["bound"]):
            assert r["evals"] == case["evals"], (case["evals"], r["evals"])
            np.testing.assert_allclose(r["x"], case["x"], rtol=1e-8, atol=1e-10)
        else:
            np.testing.assert_allclose(r["x"], case["x"], rtol=1e-4, atol=1e-5)


def test_lbfgsb_multi_and_em():
    import pllharness as PH
    fns, ns = [0, 3, 1], [4, 6, 3]
    x0 = [1.2, 1, 1.2, 1] + [1] * 6 + [1] * 3
    lo, hi, bd = [0.001] * 13, [5] * 4 + [1000] * 6 + [2.5] * 3, [2] * 13
    m = PH.test_lbfgsb_multi("b200", fns, ns, x0, lo, hi, bd)
    off = 0
    for fn, n in zip(fns, ns):
        s = PH.test_lbfgsb("b200", fn, x0[off:off + n], lo[off:off + n], hi[off:off + n], bd[off:off + n])
        np.testing.assert_allclose(m["x"][off:off + n], s["x"], rtol=1e-12)  # lock-step == one by one
        off += n
    if have_reference():
        r = PH.test_lbfgsb_multi("ref", fns, ns, x0, lo, hi, bd)
        np.testing.assert_allclose(m["x"], r["x"], rtol=1e-8)
        assert abs(m["score"] - r["score"]) < 1e-9 * abs(r["score"])
    rng = np.random.RandomState(OPT_GOLD["em"]["seed"])
    base = rng.rand(200, 4) * np.array([1, 2, 0.5, 3.0])
    sw = rng.randint(1, 4, 200)
    w = PH.test_em("b200", base, sw, [0.25] * 4)
    np.testing.assert_allclose(w, OPT_GOLD["em"]["w"], rtol=1e-12)


@pytest.mark.skipif(not have_reference(), reason="oracle/_ref not built")
def test_binary_tree_and_custom_blocks_cross_load(tmp_path):
    """pll-modules binary format (src/binary/pll_binary.c): a tree + custom block written by either library
    loads in the other with the same topology size, branch lengths, labels and payload."""
    import os
    from pllharness import binary_tree_dump, binary_tree_load
    for tips, seed in ((5, 1), (40, 7), (333, 3)):
        got = {}
        for writer in ("ref", "b200"):
            path = tmp_path / ("tree_%s_%d.bin" % (writer, tips))
            binary_tree_dump(writer, path, tips, seed)
            for reader in ("ref", "b200"):
                got[writer, reader] = binary_tree_load(reader, path)
        sizes = {w: os.path.getsize(tmp_path / ("tree_%s_%d.bin" % (w, tips))) for w in ("ref", "b200")}
        assert sizes["ref"] == sizes["b200"]
        base = got["ref", "ref"]
        assert base["tips"] == tips and base["blocks"] == 2 and base["label_sum"] > 0
        for k, v in got.items():
            assert v["tips"] == tips and v["blocks"] == 2, k
            assert abs(v["length"] - base["length"]) < 1e-12, k
            assert v["label_sum"] == base["label_sum"], k
            assert v["custom_sum"] == base["custom_sum"], k
