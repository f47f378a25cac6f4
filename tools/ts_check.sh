#!/bin/bash
# builder tool: stage-by-stage checks of the two-stage tridiagonalisation, each size in its own process under a timeout
cd "$(dirname "$0")/.."
out=gpurun_out/ts_check.log
: > $out
for spec in ${TS_SPECS:-"130 --model --full --q" "257 --model --full --q" "600 --model --full --q" "1024 --full --time --q" "1537 --full --time" "2560 --full --time" "4608 --full --time" "8193 --full --time"}; do
  echo "=== $spec" >> $out
  timeout -s KILL 90 python tools/ts_check.py --n $spec >> $out 2>&1
  echo "rc=$?" >> $out
done
tail -100 $out
