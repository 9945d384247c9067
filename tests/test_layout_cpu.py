"""Drop-in boundary: the public structs of include/ are byte-identical to the reference's.
tests/layout/layout_probe.c prints sizeof / offsetof of every public struct; it is compiled against include/
and compared with (a) the committed output for the reference headers, (b) the reference headers themselves
where /root/reference exists.  `python tests/test_layout_cpu.py --regen` rewrites the golden from (b)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROBE = os.path.join(ROOT, "tests", "layout", "layout_probe.c")
GOLD = os.path.join(ROOT, "tests", "golden", "struct_layout.json")
REF = "/root/reference"
REF_INC = [REF + "/libs/libpll/src", REF + "/src/tree", REF + "/src/optimize", REF + "/src/algorithm",
           REF + "/src/binary", REF + "/src/util", REF + "/src/msa", REF + "/src"]


def probe(includes, tmp):
    exe = os.path.join(str(tmp), "probe")
    subprocess.run(["gcc"] + ["-I" + i for i in includes] + [PROBE, "-o", exe], check=True)
    out = subprocess.run([exe], check=True, stdout=subprocess.PIPE, text=True).stdout
    return [l.strip() for l in out.splitlines() if l.strip()]


def test_public_structs_match_the_committed_reference_layout(tmp_path):
    mine = probe([os.path.join(ROOT, "include")], tmp_path)
    gold = json.load(open(GOLD))["layout"]
    assert len(mine) > 100
    assert mine == gold


@pytest.mark.skipif(not os.path.isdir(REF + "/libs/libpll/src"), reason="reference headers not present")
def test_public_structs_match_the_reference_headers(tmp_path):
    mine = probe([os.path.join(ROOT, "include")], tmp_path)
    (tmp_path / "r").mkdir()
    ref = probe(REF_INC, tmp_path / "r")
    assert mine == ref
    assert ref == json.load(open(GOLD))["layout"], "tests/golden/struct_layout.json is stale: --regen"


if __name__ == "__main__" and "--regen" in sys.argv:
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        json.dump({"generated_by": "tests/test_layout_cpu.py --regen (reference headers of /root/reference)",
                   "layout": probe(REF_INC, d)}, open(GOLD, "w"), indent=0)
