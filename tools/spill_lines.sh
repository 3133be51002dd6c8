#!/bin/bash
# usage: tools/spill_lines.sh <object.o> <function-substring>   -> source lines that carry STL/LDL (spills)
set -e
T=$(mktemp -d); cd $T
cuobjdump -xelf all "$1" >/dev/null
nvdisasm -g -c *.cubin | awk -v fn="$2" '
  /^\s*\.text\./ {infn = index($0, fn) > 0}
  /\/\/## File/ {f=$0}
  infn && /STL|LDL/ {cnt[f]++}
  END {for (k in cnt) print cnt[k], k}' | sort -rn | head -${3:-25}
rm -rf $T
