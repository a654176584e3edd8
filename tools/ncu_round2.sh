#!/bin/bash
# tools/ncu_round2.sh [tag] -- the ncu half of profile_round2.sh alone: launch list of a short C3 run, `--set full` of one Newton step's
# solve and frontal launches, and the k_batch_solve capture (each only after its plain command exited 0)
TAG=${1:-r2g}
O=gpurun_out
mkdir -p $O
cd "$(dirname "$0")/.."
python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 12 --no-cpu-baseline > $O/${TAG}_plain_c3.log 2>&1 || { echo "plain c3 failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file $O/${TAG}_launches_c3.csv \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 12 --no-cpu-baseline > $O/${TAG}_ncu_c3.log 2>&1; echo "launch list rc=$?"
read NSOLVE NFRONT <<< $(python - <<PY
import csv
rows=list(csv.reader(open("$O/${TAG}_launches_c3.csv")))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]; kn=rows[hi].index('Kernel Name')
names=[r[kn] for r in rows[hi+1:] if len(r)>kn]
ns=sum(1 for x in names if 'k_fwd' in x or 'k_bwd' in x or 'k_single_fwd' in x)
nf=sum(1 for x in names if 'k_big_panel' in x or 'k_big_assemble' in x or 'k_front_smem' in x)
print(ns//24, nf//24)
PY
)
echo "solve launches per Newton step: $NSOLVE, frontal launches per Newton step: $NFRONT"
ncu --set full --clock-control none -k 'regex:k_fwd|k_bwd' -s $NSOLVE -c $NSOLVE -f -o $O/${TAG}_c3_solve \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 3 --no-cpu-baseline > $O/${TAG}_ncu_c3solve.log 2>&1; echo "solve rc=$?"
[ -f $O/${TAG}_c3_solve.ncu-rep ] && ncu -i $O/${TAG}_c3_solve.ncu-rep --page raw --csv > $O/${TAG}_c3_solve.csv 2>/dev/null && rm -f $O/${TAG}_c3_solve.ncu-rep
ncu --set full --clock-control none -k 'regex:k_big_panel|k_big_assemble|k_front_smem' -s $NFRONT -c $NFRONT -f -o $O/${TAG}_c3_front \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 3 --no-cpu-baseline > $O/${TAG}_ncu_c3front.log 2>&1; echo "front rc=$?"
[ -f $O/${TAG}_c3_front.ncu-rep ] && ncu -i $O/${TAG}_c3_front.ncu-rep --page raw --csv > $O/${TAG}_c3_front.csv 2>/dev/null && rm -f $O/${TAG}_c3_front.ncu-rep
bash tools/ncu_c5.sh $TAG > $O/${TAG}_ncu_c5_steps.log 2>&1; tail -2 $O/${TAG}_ncu_c5_steps.log
