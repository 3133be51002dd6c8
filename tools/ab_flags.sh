# what-if timings of the tcgen05 forward through its debug flags (wrong results by construction; timing only)
for f in ${FLAGS:-0 8 32 40 0}; do echo "flags=$f"; RQ_TC_DBG_FLAGS=$f python tools/ab_fwd.py 2>&1 | grep fwd; done
