#!/bin/bash
# developer sweep of scheduling knobs on the LIVE step (graph replay, programmatic launches): us per step of cfg3, 20-step calls
run() { echo -n "$* : "; env "$@" python scripts/dev/cfg5_short.py 1 10 2>&1 | grep "   20 steps" | tail -1; }
run TG_MOM_RUN=16
run TG_MOM_RUN=8
run TG_MOM_RUN=4
run TG_MOM_RUN=2
run TG_MOM_RUN=4 TG_HEAVY_SUB=2 TG_HYBRID_GROUP=2
run TG_MOM_RUN=4 TG_HEAVY_SUB=2 TG_HYBRID_GROUP=3
run TG_MOM_RUN=4 TG_HEAVY_SUB=4 TG_HYBRID_GROUP=1
run TG_MOM_RUN=8 TG_HEAVY_SUB=4 TG_HYBRID_GROUP=3
run TG_MOM_RUN=4 TG_HEAVY_SUB=2 TG_HYBRID_GROUP=5
run TG_NO_SIDE_STREAM=1
run TG_FF_PPC=1
run TG_FF_PPC=3
