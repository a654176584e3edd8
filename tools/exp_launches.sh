#!/bin/bash
# tools/exp_launches.sh [problem] -- per-launch durations of factor + solve on one problem (factor_check under ncu, launch list only)
P=${1:-cont300}
O=gpurun_out
mkdir -p $O
python tools/factor_check.py $P > $O/fc_$P.log 2>&1 || { tail -3 $O/fc_$P.log; exit 1; }
cat $O/fc_$P.log
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file $O/launches_fc_$P.csv python tools/factor_check.py $P > $O/ncu_fc_$P.log 2>&1
echo "ncu rc=$?"
