#!/bin/bash
shape=$1; shift
for env in "$@"; do
  echo "== predict $env"
  env $env timeout 120 python scripts/run_assign.py $shape 3 predict | cut -c1-150
done
