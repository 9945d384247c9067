"""ctypes front-end of oracle/harness.c — TEST INFRASTRUCTURE.

The same C harness is built twice (against the unmodified reference -> oracle/_ref/libharness_ref.so,
and against the product -> pll-modules_b200/lib/libharness_b200.so); `Harness(kind)` drives either
through identical calls so parity tests compare like with like.
"""
import ctypes as C
import importlib.util
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "libharness_ref.so")
B200_HARNESS = os.path.join(ROOT, "pll-modules_b200", "lib", "libharness_b200.so")

ATTR_AVX2 = 1 << 2
ATTR_PATTERN_TIP = 1 << 4
ATTR_RATE_SCALERS = 1 << 9

BLO_NEWTON_FAST, BLO_NEWTON_SAFE, BLO_NEWTON_OLDFAST, BLO_NEWTON_OLDSAFE = 0, 1, 10, 11

GTR_RATES = [1.2, 2.5, 0.8, 1.1, 3.0, 1.0]
GTR_FREQS = [0.17, 0.19, 0.25, 0.39]

REDUCE_CB = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_double), C.c_size_t, C.c_int)


def load_product_package():
    spec = importlib.util.spec_from_file_location(
        "pll_modules_b200", os.path.join(ROOT, "pll-modules_b200", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def have_reference():
    return os.path.exists(REF_HARNESS)


def synthetic_model(states, seed=7):
    """A reversible model for arbitrary state counts (exchangeabilities + frequencies)."""
    if states == 4:
        return list(GTR_RATES), list(GTR_FREQS)
    if states == 20:
        import json
        with open(os.path.join(ROOT, "tests", "golden", "lg_model.json")) as f:
            m = json.load(f)
        return m["rates"], m["freqs"]
    rng = np.random.RandomState(seed + states)
    rates = (0.2 + rng.rand(states * (states - 1) // 2) * 3.0).tolist()
    rates[-1] = 1.0
    f = 0.5 + rng.rand(states)
    return rates, (f / f.sum()).tolist()


_libs = {}


def _lib(kind):
    if kind not in _libs:
        path = REF_HARNESS if kind == "ref" else B200_HARNESS
        lib = C.CDLL(path, mode=C.RTLD_LOCAL)
        lib.h_create.restype = C.c_void_p
        lib.h_create.argtypes = [C.c_uint, C.c_uint, C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_uint,
                                 C.c_uint, C.c_ulonglong, C.c_double, C.c_double, C.POINTER(C.c_double),
                                 C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                 C.c_int, C.c_uint, C.c_uint]
        lib.h_destroy.argtypes = [C.c_void_p]
        lib.h_compute_loglh.restype = C.c_double
        lib.h_compute_loglh.argtypes = [C.c_void_p, C.c_int]
        lib.h_compute_loglh_persite.restype = C.c_double
        lib.h_compute_loglh_persite.argtypes = [C.c_void_p, C.c_void_p]
        lib.h_partition_loglh.argtypes = [C.c_void_p, C.c_void_p]
        lib.h_local_sites.restype = C.c_uint
        lib.h_local_sites.argtypes = [C.c_void_p, C.c_uint]
        lib.h_get_clv.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p]
        lib.h_get_scaler.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p]
        lib.h_get_pmatrix.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p]
        lib.h_get_rates.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        lib.h_get_branch_lengths.argtypes = [C.c_void_p, C.c_void_p]
        lib.h_set_branch_lengths.argtypes = [C.c_void_p, C.c_void_p]
        lib.h_root_node_index.restype = C.c_uint
        lib.h_root_node_index.argtypes = [C.c_void_p]
        lib.h_set_root.argtypes = [C.c_void_p, C.c_uint]
        lib.h_node_info.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        lib.h_time_update_partials.restype = C.c_double
        lib.h_time_update_partials.argtypes = [C.c_void_p, C.c_int]
        lib.h_queue_update_partials.argtypes = [C.c_void_p]
        lib.h_time_edge_loglh.restype = C.c_double
        lib.h_time_edge_loglh.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_int]
        lib.h_time_derivatives.restype = C.c_double
        lib.h_time_derivatives.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_int]
        lib.h_update_partials_only.argtypes = [C.c_void_p, C.c_int]
        lib.h_edge_loglh.restype = C.c_double
        lib.h_edge_loglh.argtypes = [C.c_void_p, C.c_uint, C.c_uint]
        lib.h_root_loglh.restype = C.c_double
        lib.h_root_loglh.argtypes = [C.c_void_p, C.c_uint, C.c_uint]
        lib.h_derivatives.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.h_get_sumtable.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        lib.h_optimize_branch_lengths.restype = C.c_double
        lib.h_optimize_branch_lengths.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double,
                                                  C.c_double, C.c_double]
        lib.h_optimize_branch_lengths_single.restype = C.c_double
        lib.h_optimize_branch_lengths_single.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        lib.h_set_alpha.argtypes = [C.c_void_p, C.c_uint, C.c_double]
        lib.h_optimize_branch_lengths_iterative.restype = C.c_double
        lib.h_optimize_branch_lengths_iterative.argtypes = [C.c_void_p, C.c_int, C.c_double]
        lib.h_edge_loglh_multi.restype = C.c_double
        lib.h_edge_loglh_multi.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        lib.h_derivative_func.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.h_run_concurrently.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        lib.h_gamma_cats.argtypes = [C.c_double, C.c_uint, C.c_int, C.c_void_p]
        lib.h_perturb_edge.argtypes = [C.c_void_p, C.c_uint, C.c_double]
        lib.h_errmsg.restype = C.c_char_p
        lib.h_partition.restype = C.c_void_p
        lib.h_partition.argtypes = [C.c_void_p, C.c_uint]
        lib.h_treeinfo.restype = C.c_void_p
        lib.h_treeinfo.argtypes = [C.c_void_p]
        lib.h_set_reduce_cb.argtypes = [C.c_void_p, C.c_void_p, REDUCE_CB]
        lib.h_op_type_counts.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p]
        lib.h_get_operations.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        lib.h_ops_counter.restype = C.c_uint
        lib.h_ops_counter.argtypes = [C.c_void_p]
        lib.h_test_brent.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_void_p]
        lib.h_test_brent_multi.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_double, C.c_int, C.c_void_p]
        lib.h_test_lbfgsb.restype = C.c_double
        lib.h_test_lbfgsb.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_double, C.c_double, C.c_void_p]
        lib.h_test_lbfgsb_multi.restype = C.c_double
        lib.h_test_lbfgsb_multi.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_void_p]
        lib.h_test_em.argtypes = [C.c_uint, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.h_optimize_params.restype = C.c_double
        lib.h_optimize_params.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_double, C.c_double,
                                          C.c_double, C.c_void_p, C.c_int]
        lib.h_get_model.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        lib.h_algo_opt_treeinfo.restype = C.c_double
        lib.h_algo_opt_treeinfo.argtypes = [C.c_void_p, C.c_int, C.c_double]
        lib.h_get_treeinfo_params.argtypes = [C.c_void_p, C.c_void_p]
        lib.h_spr_round.restype = C.c_double
        lib.h_spr_round.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_double,
                                    C.c_double, C.c_int, C.c_double, C.c_int, C.c_double, C.c_double, C.c_int,
                                    C.c_void_p]
        lib.h_topology_signature.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.h_ancestral.restype = C.c_uint
        lib.h_ancestral.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.h_binary_dump.argtypes = [C.c_void_p, C.c_char_p]
        lib.h_binary_load.argtypes = [C.c_char_p, C.c_void_p]
        lib.h_binary_tree_dump.argtypes = [C.c_char_p, C.c_uint, C.c_uint]
        lib.h_binary_tree_load.argtypes = [C.c_char_p, C.c_void_p]
        _libs[kind] = lib
    return _libs[kind]


# ---- scalar optimisers on analytic objectives (pure host code: no device involved) ----
def test_brent(kind, fn, xmin, xguess, xmax, xtol):
    out = np.zeros(4)
    _lib(kind).h_test_brent(fn, xmin, xguess, xmax, xtol, out.ctypes.data)
    return dict(x=out[0], fx=out[1], f2x=out[2], evals=int(out[3]))


def test_brent_multi(kind, fns, mask, xmin, xguess, xmax, xtol, global_range):
    n = len(fns)
    i32 = lambda v: np.ascontiguousarray(v, dtype=np.int32)
    f64 = lambda v: np.ascontiguousarray(v, dtype=np.float64)
    fns, lo, g, hi = i32(fns), f64(xmin), f64(xguess).copy(), f64(xmax)
    m = i32(mask) if mask is not None else None
    out = np.zeros(2 * n + 1)
    rc = _lib(kind).h_test_brent_multi(n, fns.ctypes.data, m.ctypes.data if m is not None else None,
                                       lo.ctypes.data, g.ctypes.data, hi.ctypes.data, xtol, global_range,
                                       out.ctypes.data)
    return dict(rc=rc, x=out[:n].copy(), fx=out[n:2 * n].copy(), evals=int(out[2 * n]))


def test_lbfgsb(kind, fn, x0, lo, hi, bound, factr=1e7, pgtol=1e-5):
    x = np.ascontiguousarray(x0, dtype=np.float64).copy()
    lo = np.ascontiguousarray(lo, dtype=np.float64)
    hi = np.ascontiguousarray(hi, dtype=np.float64)
    b = np.ascontiguousarray(bound, dtype=np.int32)
    ev = C.c_int(0)
    score = _lib(kind).h_test_lbfgsb(fn, len(x), x.ctypes.data, lo.ctypes.data, hi.ctypes.data,
                                     b.ctypes.data, factr, pgtol, C.byref(ev))
    return dict(score=score, x=x, evals=ev.value)


def test_lbfgsb_multi(kind, fns, ns, x0, lo, hi, bound, factr=1e7, pgtol=1e-5):
    x = np.ascontiguousarray(x0, dtype=np.float64).copy()
    lo = np.ascontiguousarray(lo, dtype=np.float64)
    hi = np.ascontiguousarray(hi, dtype=np.float64)
    b = np.ascontiguousarray(bound, dtype=np.int32)
    fns = np.ascontiguousarray(fns, dtype=np.int32)
    ns = np.ascontiguousarray(ns, dtype=np.int32)
    ev = C.c_int(0)
    score = _lib(kind).h_test_lbfgsb_multi(len(fns), fns.ctypes.data, ns.ctypes.data, x.ctypes.data,
                                           lo.ctypes.data, hi.ctypes.data, b.ctypes.data, factr, pgtol,
                                           C.byref(ev))
    return dict(score=score, x=x, evals=ev.value)


def test_em(kind, base, site_w, w0):
    base = np.ascontiguousarray(base, dtype=np.float64)
    sw = np.ascontiguousarray(site_w, dtype=np.uint32)
    w = np.ascontiguousarray(w0, dtype=np.float64).copy()
    _lib(kind).h_test_em(base.shape[0], base.shape[1], base.ctypes.data, sw.ctypes.data, w.ctypes.data)
    return w


def binary_tree_dump(kind, path, tips, seed):
    """tree + custom block into a sequential pll-modules binary file (no partition, no device)."""
    assert _lib(kind).h_binary_tree_dump(str(path).encode(), tips, seed)


def binary_tree_load(kind, path):
    out = np.zeros(5)
    assert _lib(kind).h_binary_tree_load(str(path).encode(), out.ctypes.data)
    return dict(tips=int(out[0]), length=out[1], label_sum=out[2], custom_sum=out[3], blocks=int(out[4]))


def binary_load(kind, path):
    """Load a checkpoint written by Harness.binary_dump (either library) and evaluate it as stored."""
    out = np.zeros(7)
    assert _lib(kind).h_binary_load(str(path).encode(), out.ctypes.data)
    return dict(loglh=out[0], loglh_after_clv_block=out[1], tips=int(out[2]), length=out[3],
                custom_sum=out[4], blocks=int(out[5]), attributes=int(out[6]))


class Harness:
    """One tree + partitions + treeinfo, in the reference ('ref') or the product ('b200')."""

    def __init__(self, kind, tips, sites, states=4, rate_cats=4, attributes=ATTR_PATTERN_TIP,
                 seed=42, brlen_mean=0.1, brlen_scale=1.0, alpha=0.7, pinv=0.0, linkage=0,
                 rank=0, nranks=1, models=None):
        self.kind = kind
        self.lib = _lib(kind)
        sites = list(sites) if isinstance(sites, (list, tuple)) else [sites]
        nparts = len(sites)
        states = list(states) if isinstance(states, (list, tuple)) else [states] * nparts
        alphas = list(alpha) if isinstance(alpha, (list, tuple)) else [alpha] * nparts
        pinvs = list(pinv) if isinstance(pinv, (list, tuple)) else [pinv] * nparts
        subst, freqs = [], []
        for p in range(nparts):
            r, f = models[p] if models else synthetic_model(states[p])
            subst += list(r)
            freqs += list(f)
        if kind == "ref":
            attributes |= ATTR_AVX2
        self.tips, self.nparts, self.states, self.rate_cats = tips, nparts, states, rate_cats
        self.attributes = attributes
        arr = lambda t, v: (t * len(v))(*v)
        self.h = self.lib.h_create(tips, nparts, arr(C.c_uint, states), arr(C.c_uint, sites), rate_cats,
                                   attributes, seed, brlen_mean, brlen_scale, arr(C.c_double, alphas),
                                   arr(C.c_double, pinvs), arr(C.c_double, subst), arr(C.c_double, freqs),
                                   linkage, rank, nranks)
        if not self.h:
            raise RuntimeError("h_create failed (%s): %s" % (kind, self.lib.h_errmsg().decode()))
        self.sites = [self.lib.h_local_sites(self.h, p) for p in range(nparts)]
        self._cb = None

    def close(self):
        if self.h:
            self.lib.h_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # --- hot path ---
    def loglh(self, incremental=0):
        return self.lib.h_compute_loglh(self.h, incremental)

    def partition_loglh(self):
        out = np.zeros(self.nparts)
        self.lib.h_partition_loglh(self.h, out.ctypes.data)
        return out

    def loglh_persite(self):
        out = np.zeros(self.sites[0])
        v = self.lib.h_compute_loglh_persite(self.h, out.ctypes.data)
        return v, out

    def clv(self, clv_index, p=0):
        out = np.zeros((self.sites[p], self.rate_cats, self.states[p]))
        assert self.lib.h_get_clv(self.h, p, clv_index, out.ctypes.data)
        return out

    def scaler(self, idx, p=0):
        n = self.sites[p] * (self.rate_cats if self.attributes & ATTR_RATE_SCALERS else 1)
        out = np.zeros(n, dtype=np.uint32)
        assert self.lib.h_get_scaler(self.h, p, idx, out.ctypes.data)
        return out

    def pmatrix(self, idx, p=0):
        out = np.zeros((self.rate_cats, self.states[p], self.states[p]))
        assert self.lib.h_get_pmatrix(self.h, p, idx, out.ctypes.data)
        return out

    def rates(self, p=0):
        out = np.zeros(self.rate_cats)
        self.lib.h_get_rates(self.h, p, out.ctypes.data)
        return out

    def branch_lengths(self):
        out = np.zeros(2 * self.tips - 3)
        self.lib.h_get_branch_lengths(self.h, out.ctypes.data)
        return out

    def set_branch_lengths(self, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        self.lib.h_set_branch_lengths(self.h, v.ctypes.data)

    def root(self):
        return self.lib.h_root_node_index(self.h)

    def set_root(self, node_index):
        assert self.lib.h_set_root(self.h, node_index)

    def node_info(self, node_index):
        out = np.zeros(5, dtype=np.int32)
        self.lib.h_node_info(self.h, node_index, out.ctypes.data)
        return dict(clv=int(out[0]), scaler=int(out[1]), pmatrix=int(out[2]), back=int(out[3]), tip=bool(out[4]))

    def edge_loglh(self, node_index, p=0):
        return self.lib.h_edge_loglh(self.h, p, node_index)

    def root_loglh(self, node_index, p=0):
        return self.lib.h_root_loglh(self.h, p, node_index)

    def derivatives(self, node_index, brlens, p=0):
        b = np.ascontiguousarray(brlens, dtype=np.float64)
        df, ddf = np.zeros(len(b)), np.zeros(len(b))
        ok = self.lib.h_derivatives(self.h, p, node_index, b.ctypes.data, len(b), df.ctypes.data, ddf.ctypes.data)
        if not ok:
            raise RuntimeError(self.lib.h_errmsg().decode())
        return df, ddf

    def sumtable(self, p=0):
        out = np.zeros((self.sites[p], self.rate_cats, self.states[p]))
        assert self.lib.h_get_sumtable(self.h, p, out.ctypes.data)
        return out

    def optimize_branch_lengths(self, method=BLO_NEWTON_FAST, radius=-1, max_iters=4, lh_epsilon=0.1,
                                brlen_min=1e-4, brlen_max=100.0):
        return self.lib.h_optimize_branch_lengths(self.h, method, radius, max_iters, lh_epsilon,
                                                  brlen_min, brlen_max)

    def optimize_branch_lengths_single(self, radius=-1, smoothings=2, tolerance=0.1):
        return self.lib.h_optimize_branch_lengths_single(self.h, radius, smoothings, tolerance)

    def optimize_params(self, which, p=0, factr=1e7, pgtol=1e-4, alpha_guess=0.7, symmetries=None,
                        onedim=False):
        sym = np.ascontiguousarray(symmetries, dtype=np.int32) if symmetries is not None else None
        return self.lib.h_optimize_params(self.h, p, which, factr, pgtol, alpha_guess,
                                          sym.ctypes.data if sym is not None else None, int(onedim))

    def model(self, p=0):
        K, s = self.rate_cats, self.states[p]
        nr = s * (s - 1) // 2
        out = np.zeros(2 * K + 1 + nr + s)
        self.lib.h_get_model(self.h, p, out.ctypes.data)
        return dict(rates=out[:K].copy(), weights=out[K:2 * K].copy(), pinv=out[2 * K],
                    subst=out[2 * K + 1:2 * K + 1 + nr].copy(), freqs=out[2 * K + 1 + nr:].copy())

    def spr_round(self, radius_min=1, radius_max=5, ntopol_keep=5, thorough=False, blo_method=BLO_NEWTON_FAST,
                  bl_min=1e-6, bl_max=100.0, smoothings=32, epsilon=0.1, cutoff=True, subtree_cutoff=1.0,
                  lh_eps_triplet=1000.0, fast_clv_updates=True):
        """pllmod_algo_spr_round; returns (logL, cutoff stats)."""
        co = np.zeros(3)
        lh = self.lib.h_spr_round(self.h, radius_min, radius_max, ntopol_keep, int(thorough), blo_method, bl_min,
                                  bl_max, smoothings, epsilon, int(cutoff), subtree_cutoff, lh_eps_triplet,
                                  int(fast_clv_updates), co.ctypes.data)
        return lh, co

    def topology_signature(self):
        n = 2 * self.tips - 3
        hs, ln = np.zeros(n, dtype=np.uint64), np.zeros(n)
        self.lib.h_topology_signature(self.h, hs.ctypes.data, ln.ctypes.data)
        return hs, ln

    ALGO_OPT = dict(alpha=0, pinv=1, subst_rates=2, frequencies=3, alpha_pinv=4, rates_weights=5,
                    brlen_scalers=6, brlen=7)

    def algo_opt_treeinfo(self, what, tolerance=1e-3):
        """pllmod_algo_opt_*_treeinfo over all partitions; returns -logL."""
        return self.lib.h_algo_opt_treeinfo(self.h, self.ALGO_OPT[what], tolerance)

    def ancestral(self):
        """pllmod_treeinfo_compute_ancestral: (inner nodes by clv_index, concatenated [site][state] per partition)"""
        span = sum(self.sites[p] * self.states[p] for p in range(self.nparts))
        out = np.zeros((self.tips - 2, span))
        lab = C.c_double(0.0)
        n = self.lib.h_ancestral(self.h, out.ctypes.data, C.byref(lab))
        assert n == self.tips - 2, self.lib.h_errmsg().decode()
        return out, lab.value

    def binary_dump(self, path):
        """partition 0 + tree + branch lengths + root CLV as blocks 10..13 of a random-access file"""
        assert self.lib.h_binary_dump(self.h, str(path).encode())

    def treeinfo_params(self):
        out = np.zeros(2 * self.nparts)
        self.lib.h_get_treeinfo_params(self.h, out.ctypes.data)
        return dict(alphas=out[:self.nparts].copy(), brlen_scalers=out[self.nparts:].copy())

    def set_alpha(self, alpha, p=0):
        assert self.lib.h_set_alpha(self.h, p, alpha)

    def perturb_edge(self, node_index, new_length):
        self.lib.h_perturb_edge(self.h, node_index, new_length)

    def time_update_partials(self, reps=1):
        return self.lib.h_time_update_partials(self.h, reps)

    def time_edge_loglh(self, node_index, reps=10, p=0):
        return self.lib.h_time_edge_loglh(self.h, p, node_index, reps)

    def time_derivatives(self, node_index, reps=10, p=0):
        return self.lib.h_time_derivatives(self.h, p, node_index, reps)

    def queue_update_partials(self):
        self.lib.h_queue_update_partials(self.h)

    def update_partials_only(self, incremental=0):
        return self.lib.h_update_partials_only(self.h, incremental)

    def partition_ptr(self, p=0):
        return self.lib.h_partition(self.h, p)

    def set_reduce_cb(self, fn):
        """fn(numpy view of doubles, op) performs an in-place cross-rank reduction."""
        def _cb(ctx, data, n, op):
            view = np.ctypeslib.as_array(data, shape=(n,))
            fn(view, op)
        self._cb = REDUCE_CB(_cb)
        self.lib.h_set_reduce_cb(self.h, None, self._cb)

    def set_reduce_native(self, pkg, rank, world, allgather_bytes):
        """Connect partition 0's CUDA-IPC exchange over `world` ranks (allgather_bytes(bytes) -> list of every
        rank's bytes) and make the library's own collective, pllb200_exchange_reduce_cb, the treeinfo's
        parallel_reduce_cb: logL and derivative sums then cross NVLink inside the library."""
        part = self.partition_ptr(0)
        every = allgather_bytes(pkg.exchange_create(part))
        pkg.exchange_connect(part, rank, world, b"".join(every))
        fn = C.cast(pkg.load().pllb200_exchange_reduce_cb, C.c_void_p).value
        self._cb = REDUCE_CB(fn)
        self.lib.h_set_reduce_cb(self.h, C.c_void_p(part), self._cb)

    def optimize_branch_lengths_iterative(self, smoothings=2, tolerance=1e-3):
        return self.lib.h_optimize_branch_lengths_iterative(self.h, smoothings, tolerance)

    def edge_loglh_multi(self, node_index, persite=False):
        buf = np.zeros(sum(self.sites)) if persite else None
        ll = self.lib.h_edge_loglh_multi(self.h, node_index, buf.ctypes.data if persite else None)
        return (ll, buf) if persite else ll

    def derivative_func(self, node_index, brlens, p=0):
        b = np.ascontiguousarray(brlens, dtype=np.float64)
        df, ddf = np.zeros(len(b)), np.zeros(len(b))
        assert self.lib.h_derivative_func(self.h, p, node_index, b.ctypes.data, len(b), df.ctypes.data, ddf.ctypes.data)
        return df, ddf

    def op_type_counts(self, p=0):
        out = np.zeros(3, dtype=np.uint32)
        self.lib.h_op_type_counts(self.h, p, self.tips - 2, out.ctypes.data)
        return dict(tt=int(out[0]), ti=int(out[1]), ii=int(out[2]))

    def operations(self, nops):
        """pll_operation_t rows of the last traversal (after loglh(0): tips - 2 of them)."""
        out = np.zeros((nops, 8), dtype=np.int32)
        self.lib.h_get_operations(self.h, nops, out.ctypes.data)
        return out

    def ops_counter(self):
        return self.lib.h_ops_counter(self.h)


def run_concurrently(harnesses, threaded=True):
    """Every harness (same kind) from its own pthread at once: [logL, -logL after one BLO call, logL] each."""
    lib = harnesses[0].lib
    arr = (C.c_void_p * len(harnesses))(*[h.h for h in harnesses])
    out = np.zeros(3 * len(harnesses))
    assert lib.h_run_concurrently(arr, len(harnesses), 1 if threaded else 0, out.ctypes.data)
    return out.reshape(len(harnesses), 3)


def gamma_cats(kind, alpha, cats, mode=0):
    out = np.zeros(cats)
    assert _lib(kind).h_gamma_cats(alpha, cats, mode, out.ctypes.data)
    return out
