"""One full traversal at 61 states (for ncu captures of partial_dmma_kernel)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from pllharness import Harness
with Harness("b200", tips=24, sites=20000, states=61) as h:
    print(h.loglh(0), h.time_update_partials(2))
