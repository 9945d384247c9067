"""Debug aid: CLVs of one case computed by two builds of the library (lib/libpll_b200.<variant>), first differences.
usage: python scripts/probes/clv_diff.py <variantA> <variantB>   (run on the GPU box)"""
import os, shutil, subprocess, sys
import numpy as np
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
L = os.path.join(ROOT, "pll-modules_b200", "lib")
if len(sys.argv) == 4:   # child: compute and save
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from pllharness import Harness
    AMBIG = 1 << 40
    kw = dict(tips=80, sites=515, seed=AMBIG | 7)
    with Harness("b200", **kw) as h:
        ll = h.loglh()
        ops = h.operations(78)
        clv = np.stack([h.clv(80 + i) for i in range(78)])
    np.savez(sys.argv[3], clv=clv, ops=ops, ll=ll)
    sys.exit(0)
out = {}
for v in sys.argv[1:3]:
    shutil.copy(os.path.join(L, "libpll_b200." + v), os.path.join(L, "libpll_b200.so"))
    f = "/tmp/clv_%s.npz" % v
    subprocess.check_call([sys.executable, __file__, v, "child", f])
    out[v] = np.load(f)
a, b = out[sys.argv[1]], out[sys.argv[2]]
print("logl", float(a["ll"]), float(b["ll"]))
bad = np.argwhere((a["clv"] != b["clv"]).any(axis=(2, 3)))
print("mismatching (node,site) pairs:", len(bad), "of", a["clv"].shape[0] * a["clv"].shape[1])
ops = a["ops"]
byparent = {int(o[0]): o for o in ops}
nodes = sorted(set(int(n) for n, s in bad))
for n in nodes[:12]:
    sites = [int(s) for nn, s in bad if nn == n]
    o = byparent.get(80 + n)
    kind = "".join("T" if c < 80 else "I" for c in (o[2], o[5])) if o is not None else "?"
    print("node", 80 + n, kind, "children", (int(o[2]), int(o[5])) if o is not None else None, "nbad", len(sites), "warps", sorted(set(s // 32 for s in sites))[:10], "first sites", sites[:8])
if len(bad):
    n, s = bad[0]
    print("A", a["clv"][n, s].ravel()); print("B", b["clv"][n, s].ravel())
