"""Generates tests/golden/* from the UNMODIFIED reference (oracle/_ref, built by oracle/build_ref.sh).

Run in the build container (needs /root/reference to have been compiled into oracle/_ref):
    python oracle/gen_golden.py
Outputs (committed, small):
    tests/golden/lg_model.json      LG exchangeabilities + frequencies read from the reference library's
                                    exported tables pll_aa_rates_lg / pll_aa_freqs_lg (maps.c:288,1472)
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


def main():
    os.makedirs(GOLD, exist_ok=True)
    dump_lg()
    from pllharness import Harness
    res = {}
    for name, kw in CASES:
        res[name] = {"args": kw, "ref": run_case(Harness, kw)}
        print(name, res[name]["ref"]["loglh"])
    with open(os.path.join(GOLD, "ref_cases.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
