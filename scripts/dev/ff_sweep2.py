import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ff_sweep import run
run(10, {"TG_FARFIELD": "0"})
for hs, pct, hg in (("8", "0", "1"), ("8", "100", "2"), ("4", "0", "1"), ("4", "100", "2"), ("4", "100", "3"), ("2", "100", "2"), ("2", "70", "3")):
    run(10, {"TG_HEAVY_SUB": hs, "TG_HYBRID_PCT": pct, "TG_HYBRID_GROUP": hg})
run(80, {"TG_FARFIELD": "0"})
for hs, hg in (("8", "2"), ("8", "3"), ("4", "3")):
    run(80, {"TG_HEAVY_SUB": hs, "TG_HYBRID_PCT": "100", "TG_HYBRID_GROUP": hg})
