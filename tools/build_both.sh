#!/bin/bash
# builds the release library and an -DRQ_EXPERIMENTS copy (librqengine_exp.so) next to it
set -e
cd "$(dirname "$0")/.."
python -m rq_vae_recommender_b200.build --experiments --force >/dev/null
cp rq_vae_recommender_b200/librqengine.so rq_vae_recommender_b200/librqengine_exp.so
python -m rq_vae_recommender_b200.build --force >/dev/null
echo built
