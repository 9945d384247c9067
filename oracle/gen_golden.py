"""Generates tests/golden/* from the UNMODIFIED reference (oracle/_ref, built by oracle/build_ref.sh).

Run in the build container (needs /root/reference to have been compiled into oracle/_ref):
    python oracle/gen_golden.py
Outputs (committed, small):
    tests/golden/lg_model.json      LG exchangeabilities + frequencies read from the reference library's
                                    exported tables pll_aa_rates_lg / pll_aa_freqs_lg (maps.c:288,1472)
    tests/golden/ref_opt.json       Brent / L-BFGS-B / EM on analytic objectives and model-parameter optimisation
                                    (pllmod_opt_optimize_onedim / _multidim) results of the reference
    tests/golden/ref_cases.json     logL / derivatives / scaler statistics / BLO results of the reference
                                    AVX2 path on the seeded synthetic inputs of oracle/harness.c
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLD = os.path.join(ROOT, "tests", "golden")


def dump_lg():
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libpll_ref.so"), mode=C.RTLD_LOCAL)
    rates = (C.c_double * 190).in_dll(ref, "pll_aa_rates_lg")
    freqs = (C.c_double * 20).in_dll(ref, "pll_aa_freqs_lg")
    with open(os.path.join(GOLD, "lg_model.json"), "w") as f:
        json.dump({"rates": list(rates), "freqs": list(freqs)}, f)


CASES = [
    # name, kwargs for Harness
    ("dna16", dict(tips=16, sites=1000, states=4, rate_cats=4)),
    ("dna16_pinv", dict(tips=16, sites=1000, states=4, rate_cats=4, pinv=0.2)),
    ("dna16_k1", dict(tips=16, sites=777, states=4, rate_cats=1)),
    ("dna16_notip", dict(tips=16, sites=500, states=4, rate_cats=4, attributes=0)),
    ("dna200_deep", dict(tips=200, sites=300, states=4, rate_cats=4, brlen_scale=8.0)),
    ("dna200_deep_ratescalers", dict(tips=200, sites=300, states=4, rate_cats=4, brlen_scale=8.0,
                                     attributes=(1 << 4) | (1 << 9))),
    ("aa12", dict(tips=12, sites=200, states=20, rate_cats=4)),
    ("odd5", dict(tips=10, sites=300, states=5, rate_cats=3)),
    # ascertainment bias (attributes: PATTERN_TIP 1<<4, AB_FLAG 1<<8, type 1..3 << 5)
    ("dna12_asc_lewis", dict(tips=12, sites=300, states=4, rate_cats=4, attributes=(1 << 4) | (1 << 8) | (1 << 5))),
    ("dna12_asc_felsenstein_notip", dict(tips=12, sites=300, states=4, rate_cats=4, attributes=(1 << 8) | (2 << 5))),
    ("aa12_asc_stamatakis", dict(tips=12, sites=150, states=20, rate_cats=4, attributes=(1 << 4) | (1 << 8) | (3 << 5))),
    ("aa12_asc_lewis", dict(tips=12, sites=150, states=20, rate_cats=4, attributes=(1 << 4) | (1 << 8) | (1 << 5))),
    ("dna150_deep_asc_felsenstein", dict(tips=150, sites=200, states=4, rate_cats=4, brlen_scale=8.0,
                                         attributes=(1 << 4) | (1 << 8) | (2 << 5))),
    ("multi", dict(tips=20, sites=[300, 150, 200], states=[4, 20, 4], rate_cats=4, alpha=[0.3, 0.5, 1.1])),
]


def run_case(H, kw):
    out = {}
    with H("ref", **kw) as h:
        out["loglh"] = h.loglh()
        out["partition_loglh"] = h.partition_loglh().tolist()
        r = h.root()
        brl = [1e-4, 0.01, 0.1, 0.5, 2.0]
        df, ddf = h.derivatives(r, brl)
        out["deriv_brlens"] = brl
        out["df"] = df.tolist()
        out["ddf"] = ddf.tolist()
        sc = [int(h.scaler(i).sum()) for i in range(kw["tips"] - 2)]
        out["scaler_sums"] = sc
        info = h.node_info(r)
        out["root_clv_sum"] = float(h.clv(info["clv"]).sum())
        out["rates"] = h.rates().tolist()
        out["pmatrix0"] = h.pmatrix(0).ravel().tolist()[:16]
        out["blo_fast"] = h.optimize_branch_lengths(method=0, radius=-1, max_iters=2, lh_epsilon=0.1)
        out["blo_brlens"] = h.branch_lengths().tolist()
    return out


# ---- optimiser goldens -> tests/golden/ref_opt.json ----------------------
BRENT_CASES = [(fn, lo, g, hi, tol) for fn in range(4)
               for (lo, g, hi, tol) in [(0.02, 0.5, 100.0, 1e-4), (1e-4, 2.0, 10.0, 1e-3), (0.02, 0.02, 5.0, 1e-2)]]
LBFGSB_CASES = [
    dict(fn=0, x0=[1.2, 1, 1.2, 1, 0.5], lo=[0.001] * 5, hi=[5] * 5, bound=[2] * 5),
    dict(fn=1, x0=[1] * 6, lo=[0.001] * 6, hi=[2.5] * 6, bound=[2] * 6),
    dict(fn=2, x0=[4, 3, 2, 1, 0.5], lo=[0.001] * 5, hi=[100] * 5, bound=[2, 2, 1, 1, 0]),
    dict(fn=3, x0=[1] * 6, lo=[0.001] * 6, hi=[1000] * 6, bound=[2] * 6),
    dict(fn=3, x0=[1] * 6, lo=[0.001] * 6, hi=[2.0] * 6, bound=[2] * 6),
    dict(fn=2, x0=[1.2, 1, 1.2, 1, .5, .7], lo=[0.001] * 6, hi=[5] * 6, bound=[0] * 6),
]
# PLLMOD_OPT_PARAM_*: SUBST_RATES 1, ALPHA 2, PINV 4, FREQUENCIES 8, FREE_RATES 256, RATE_WEIGHTS 512
MODEL_OPT_CASES = [
    ("dna16_alpha_brent", dict(tips=16, sites=1000), dict(which=2, onedim=True, pgtol=1e-3, alpha_guess=0.7)),
    ("dna16_pinv_brent", dict(tips=16, sites=1000, pinv=0.2), dict(which=4, onedim=True, pgtol=1e-3)),
    ("dna16_rates_lbfgsb", dict(tips=16, sites=1000), dict(which=1, factr=1e7, pgtol=1e-4)),
    ("dna16_rates_sym_lbfgsb", dict(tips=16, sites=1000),
     dict(which=1, factr=1e7, pgtol=1e-4, symmetries=[0, 1, 0, 0, 1, 0])),
    ("dna16_freqs_alpha_lbfgsb", dict(tips=16, sites=1000), dict(which=8 | 2, factr=1e7, pgtol=1e-4)),
    ("dna16_freerates_weights_lbfgsb", dict(tips=16, sites=1000), dict(which=256 | 512, factr=1e7, pgtol=1e-4)),
    ("dna16_pinv_alpha_lbfgsb", dict(tips=16, sites=1000, pinv=0.2), dict(which=4 | 2, factr=1e7, pgtol=1e-4)),
    ("aa12_alpha_brent", dict(tips=12, sites=200, states=20), dict(which=2, onedim=True, pgtol=1e-3, alpha_guess=0.7)),
]


SPR_CASES = [
    ("dna24_fast", dict(tips=24, sites=400, seed=11), dict(thorough=False, radius_max=5)),
    ("dna24_thorough", dict(tips=24, sites=400, seed=11), dict(thorough=True, radius_max=5)),
    ("dna40_fast_r10_nocutoff", dict(tips=40, sites=300, seed=5), dict(thorough=False, radius_max=10, cutoff=False,
                                                                     fast_clv_updates=False)),
    ("aa16_fast", dict(tips=16, sites=150, states=20, seed=3), dict(thorough=False, radius_max=4)),
    ("multi_unlinked_thorough", dict(tips=14, sites=[200, 150], states=[4, 4], alpha=[0.5, 1.0], linkage=2, seed=9),
     dict(thorough=True, radius_max=3, ntopol_keep=3)),
    # a tree of BASELINE configs[4]'s kind at a size the reference finishes in seconds
    ("dna200_fast_2k", dict(tips=200, sites=2000, seed=31), dict(thorough=False, radius_max=5)),
]


TWO_PART = dict(tips=14, sites=[300, 200], states=[4, 4], alpha=[0.5, 1.0], seed=21)
ALGO_OPT_CASES = [
    ("alpha_2parts", TWO_PART, "alpha"),
    ("pinv_2parts", dict(TWO_PART, pinv=[0.1, 0.2]), "pinv"),
    ("subst_rates_2parts", TWO_PART, "subst_rates"),
    ("frequencies_2parts", TWO_PART, "frequencies"),
    ("alpha_pinv_2parts", dict(TWO_PART, pinv=[0.1, 0.2]), "alpha_pinv"),
    ("rates_weights_2parts_unlinked", dict(TWO_PART, linkage=2), "rates_weights"),
    ("rates_weights_1part", dict(tips=14, sites=400, seed=21), "rates_weights"),
    ("brlen_scalers_scaled", dict(TWO_PART, linkage=1), "brlen_scalers"),
    ("brlen_2parts", TWO_PART, "brlen"),
    ("subst_rates_dna_aa", dict(tips=10, sites=[200, 80], states=[4, 20], alpha=[0.5, 1.0], seed=4), "alpha"),
]


def spr_case(PH, name, kw, skw):
    with PH.Harness("ref", **kw) as h:
        start = h.loglh()
        lh, co = h.spr_round(**skw)
        hs, ln = h.topology_signature()
        print(name, start, lh, h.ops_counter())
        return dict(args=kw, spr=skw, start_loglh=start, loglh=lh, cutoff=co.tolist(),
                    splits=[str(int(x)) for x in hs], brlens=ln.tolist(), ops=h.ops_counter())


def add_missing_spr():
    """--spr-missing: run only the SPR cases that tests/golden/ref_opt.json does not hold yet and merge them in."""
    import pllharness as PH
    path = os.path.join(GOLD, "ref_opt.json")
    out = json.load(open(path))
    for name, kw, skw in SPR_CASES:
        if name not in out["spr"]:
            out["spr"][name] = spr_case(PH, name, kw, skw)
    with open(path, "w") as f:
        json.dump(out, f, indent=1)


def dump_opt():
    import pllharness as PH
    out = {"brent": [], "lbfgsb": [], "model": {}, "spr": {}, "algo_opt": {}}
    for name, kw, what in ALGO_OPT_CASES:
        with PH.Harness("ref", **kw) as h:
            start = h.loglh()
            score = h.algo_opt_treeinfo(what)
            after = h.loglh()
            models = [{k: (v.tolist() if hasattr(v, "tolist") else float(v)) for k, v in h.model(p).items()}
                      for p in range(h.nparts)]
            tp = {k: v.tolist() for k, v in h.treeinfo_params().items()}
            out["algo_opt"][name] = dict(args=kw, what=what, start_loglh=start, score=score, loglh_after=after,
                                         models=models, treeinfo=tp, brlens=h.branch_lengths().tolist())
            print(name, start, score, after)
    for name, kw, skw in SPR_CASES:
        out["spr"][name] = spr_case(PH, name, kw, skw)
    for fn, lo, g, hi, tol in BRENT_CASES:
        r = PH.test_brent("ref", fn, lo, g, hi, tol)
        out["brent"].append(dict(args=[fn, lo, g, hi, tol], x=float(r["x"]), fx=float(r["fx"]), evals=r["evals"]))
    for c in LBFGSB_CASES:
        r = PH.test_lbfgsb("ref", c["fn"], c["x0"], c["lo"], c["hi"], c["bound"])
        out["lbfgsb"].append(dict(args=c, score=r["score"], x=r["x"].tolist(), evals=r["evals"]))
    rng = np.random.RandomState(3)
    base = rng.rand(200, 4) * np.array([1, 2, 0.5, 3.0])
    sw = rng.randint(1, 4, 200)
    out["em"] = dict(seed=3, w=PH.test_em("ref", base, sw, [0.25] * 4).tolist())
    for name, kw, okw in MODEL_OPT_CASES:
        with PH.Harness("ref", **kw) as h:
            start = h.loglh()
            score = h.optimize_params(**okw)
            m = h.model()
            after = h.loglh()
        out["model"][name] = dict(args=kw, opt=okw, start_loglh=start, score=score, loglh_after=after,
                                  model={k: (v.tolist() if hasattr(v, "tolist") else float(v)) for k, v in m.items()})
        print(name, start, score)
    with open(os.path.join(GOLD, "ref_opt.json"), "w") as f:
        json.dump(out, f, indent=1)


def main():
    if "--spr-missing" in sys.argv:
        return add_missing_spr()
    if "--opt-only" in sys.argv:
        return dump_opt()
    os.makedirs(GOLD, exist_ok=True)
    dump_lg()
    from pllharness import Harness
    res = {}
    for name, kw in CASES:
        res[name] = {"args": kw, "ref": run_case(Harness, kw)}
        print(name, res[name]["ref"]["loglh"])
    with open(os.path.join(GOLD, "ref_cases.json"), "w") as f:
        json.dump(res, f, indent=1)
    dump_opt()


if __name__ == "__main__":
    main()
