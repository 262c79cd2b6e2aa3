import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ff_sweep import run
for hs in ("4", "2", "6"):
    for pct in ("100", "85", "70", "50", "30"):
        run(10, {"TG_HEAVY_SUB": hs, "TG_HYBRID_PCT": pct, "TG_HYBRID_GROUP": "2"})
run(10, {"TG_HEAVY_SUB": "4", "TG_HYBRID_PCT": "70", "TG_HYBRID_GROUP": "3", "TG_HEAVY_GROUP": "2"})
run(10, {"TG_HEAVY_SUB": "6", "TG_HYBRID_PCT": "70", "TG_HYBRID_GROUP": "2", "TG_HEAVY_GROUP": "1"})
