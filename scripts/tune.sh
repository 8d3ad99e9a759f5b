#!/bin/bash
# usage: tune.sh "N D K" "VAR=val VAR=val" ...   (each extra arg is one environment setting to try)
shape=$1; shift
for env in "$@"; do
  echo "== $env"
  env $env timeout 120 python scripts/run_assign.py $shape 3 | cut -c1-110
done
