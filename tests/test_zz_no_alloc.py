"""pllb200_edge_loglikelihood_noalloc (no counterpart in the reference):
the edge log-likelihood with its scratch memory in a static device array, through the C ABI, against
the existing entry point pll_compute_edge_loglikelihood.  Collected last on purpose (file name).
Tolerances as in test_gpu_parity.py: logL 1e-8 relative, per-site values 1e-9."""
import ctypes as C

import numpy as np
import pytest

from pllharness import Harness, load_product_package

pytestmark = pytest.mark.gpu
NOALLOC_MAX_SITES = 262144


def _api():
    lib = load_product_package().load()
    lib.pll_compute_edge_loglikelihood.restype = C.c_double
    lib.pll_compute_edge_loglikelihood.argtypes = [C.c_void_p, C.c_uint, C.c_int, C.c_uint, C.c_int, C.c_uint,
                                                   C.c_void_p, C.c_void_p]
    lib.pllb200_edge_loglikelihood_noalloc.restype = C.c_int
    lib.pllb200_edge_loglikelihood_noalloc.argtypes = [C.c_void_p, C.c_uint, C.c_int, C.c_uint, C.c_int, C.c_uint,
                                                       C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def _edge(h, node):
    a = h.node_info(node)
    b = h.node_info(a["back"])
    return a["clv"], a["scaler"], b["clv"], b["scaler"], a["pmatrix"]


@pytest.mark.parametrize("kw", [
    dict(tips=24, sites=5000),                          # DNA Gamma-4: logl_dna_kernel
    dict(tips=12, sites=777, states=20),                # protein: logl_aa_kernel
    dict(tips=10, sites=300, states=7, rate_cats=3),    # generic kernel
    dict(tips=30, sites=2000, brlen_scale=10.0),        # scaler counts in play
], ids=lambda kw: "-".join("%s%s" % (k[:3], v) for k, v in kw.items()))
def test_noalloc_entry_point_equals_the_existing_one_twice_in_a_row(kw):
    lib = _api()
    with Harness("b200", **kw) as h:
        h.loglh(0)                                       # CLVs, P-matrices and the model block on the device
        part = h.partition_ptr(0)
        sites = h.sites[0]
        freqs = (C.c_uint * kw.get("rate_cats", 4))()
        root = h.root()
        inner = [n for n in range(kw["tips"], kw["tips"] + 6) if not h.node_info(h.node_info(n)["back"])["tip"]]
        for node in (root, root) + tuple(inner[:1]):     # the same edge twice in a row, then another one
            pc, ps, cc, cs, m = _edge(h, node)
            if node != root:
                h.set_root(node)
                h.loglh(1)                                # orient the CLVs towards this edge
                pc, ps, cc, cs, m = _edge(h, node)
            ref_ps = np.zeros(sites)
            ref = lib.pll_compute_edge_loglikelihood(part, pc, ps, cc, cs, m, freqs, ref_ps.ctypes.data)
            for with_persite in (True, False):
                got = C.c_double(0.0)
                got_ps = np.full(sites, np.nan)
                rc = lib.pllb200_edge_loglikelihood_noalloc(part, pc, ps, cc, cs, m, freqs, C.byref(got),
                                                            got_ps.ctypes.data if with_persite else None)
                assert rc == 1, h.lib.h_errmsg()
                assert abs(got.value - ref) <= 1e-8 * abs(ref), (got.value, ref)
                if with_persite:
                    np.testing.assert_allclose(got_ps, ref_ps, rtol=1e-9, atol=0)
            # the existing entry point is not disturbed by the shared scratch
            again = lib.pll_compute_edge_loglikelihood(part, pc, ps, cc, cs, m, freqs, None)
            assert abs(again - ref) <= 1e-8 * abs(ref)


def test_noalloc_entry_point_rejects_partitions_beyond_its_static_scratch():
    lib = _api()
    with Harness("b200", tips=6, sites=NOALLOC_MAX_SITES + 1) as h:
        h.loglh(0)
        pc, ps, cc, cs, m = _edge(h, h.root())
        freqs = (C.c_uint * 4)()
        got = C.c_double(0.0)
        rc = lib.pllb200_edge_loglikelihood_noalloc(h.partition_ptr(0), pc, ps, cc, cs, m, freqs, C.byref(got), None)
        assert rc == 0
        ref = lib.pll_compute_edge_loglikelihood(h.partition_ptr(0), pc, ps, cc, cs, m, freqs, None)
        assert ref < 0.0
    with Harness("b200", tips=6, sites=NOALLOC_MAX_SITES) as h:   # exactly the maximum is served
        h.loglh(0)
        pc, ps, cc, cs, m = _edge(h, h.root())
        freqs = (C.c_uint * 4)()
        got = C.c_double(0.0)
        ref = lib.pll_compute_edge_loglikelihood(h.partition_ptr(0), pc, ps, cc, cs, m, freqs, None)
        assert lib.pllb200_edge_loglikelihood_noalloc(h.partition_ptr(0), pc, ps, cc, cs, m, freqs, C.byref(got), None) == 1
        assert abs(got.value - ref) <= 1e-8 * abs(ref)
