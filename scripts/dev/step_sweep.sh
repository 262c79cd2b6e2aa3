#!/bin/bash
# developer sweep of scheduling knobs on the LIVE step (graph replay, programmatic launches): us per step, 20-step calls (S stars x 10 chains)
S=${1:-1}
run() { echo -n "$* : "; env "$@" python scripts/dev/cfg5_short.py $S 10 2>&1 | grep "   20 steps" | tail -1; }
run TG_X=0
run TG_HEAVY_SUB=4 TG_HYBRID_GROUP=3
run TG_HEAVY_SUB=2 TG_HYBRID_GROUP=3
run TG_HEAVY_SUB=8 TG_HYBRID_GROUP=2
run TG_HEAVY_SUB=8 TG_HYBRID_GROUP=3
run TG_HEAVY_SUB=8 TG_HYBRID_GROUP=4
run TG_HEAVY_SUB=4 TG_HYBRID_GROUP=2
run TG_HEAVY_SUB=4 TG_HYBRID_GROUP=5
run TG_FF_PPC=1
run TG_FF_PPC=2
run TG_FF_PPC=4
