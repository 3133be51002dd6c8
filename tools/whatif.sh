#!/bin/bash
# timing what-ifs of the d=32 forward (results wrong by construction): needs the -DRQ_EXPERIMENTS library
for f in ${FLAGS:-0 1 2 4 8 3 6 12 15 0}; do echo -n "whatif=$f "; RQ_TC_WHATIF=$f python tools/ab_fwd.py rq_vae_recommender_b200/librqengine_exp.so 2>&1 | grep fwd | sed 's/.*exp.so//'; done
