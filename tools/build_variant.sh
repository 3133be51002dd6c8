#!/bin/bash
# usage: tools/build_variant.sh NAME "-DFLAG1 -DFLAG2"  -> rq_vae_recommender_b200/librqengine_NAME.so (release library untouched)
set -e
cd "$(dirname "$0")/.."
cp rq_vae_recommender_b200/librqengine.so /tmp/_lib_release.so
RQ_EXTRA_NVCC_FLAGS="$2" python -m rq_vae_recommender_b200.build --force >/dev/null
cp rq_vae_recommender_b200/librqengine.so rq_vae_recommender_b200/librqengine_$1.so
cp /tmp/_lib_release.so rq_vae_recommender_b200/librqengine.so
python - <<'PY'
import rq_vae_recommender_b200.build as b, os
open(os.path.join(b.OBJ, "digest.txt"), "w").write("stale")
PY
echo "built librqengine_$1.so"
