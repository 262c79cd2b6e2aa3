import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ff_sweep import run
run(10, {"TG_FARFIELD": "0"})
for hs in ("2", "4", "8"):
    for pct, hg in (("0", "1"), ("100", "2"), ("100", "3"), ("70", "3")):
        run(10, {"TG_HEAVY_SUB": hs, "TG_HYBRID_PCT": pct, "TG_HYBRID_GROUP": hg})
run(80, {"TG_FARFIELD": "0"})
for hs, hg in (("2", "3"), ("4", "2"), ("4", "3"), ("4", "5"), ("8", "3"), ("8", "5")):
    run(80, {"TG_HEAVY_SUB": hs, "TG_HYBRID_PCT": "100", "TG_HYBRID_GROUP": hg})
