#!/bin/bash
# tools/profile_round2.sh [tag] -- round-2 evidence on one B200 (under gpurun): GPU tests, smoke, the bench lines of both
# arms and of every workload, then (each only after its plain command exited 0) the ncu launch lists and captures.
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
cd "$(dirname "$0")/.."
step() { echo "== $*" | tee -a $O/${TAG}_steps.log; }
: > $O/${TAG}_steps.log
step "pytest -m gpu"
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?" | tee -a $O/${TAG}_steps.log; tail -2 $O/${TAG}_pytest_gpu.log
step "smoke"
python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > $O/${TAG}_smoke.log 2>&1; tail -1 $O/${TAG}_smoke.log
step "bench default (c5 + kkt + parity)"
python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench_c5.json 2> $O/${TAG}_bench_c5.err; echo "rc=$?"
step "bench --impl reference"
python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_c5_reference.json 2> $O/${TAG}_bench_c5_reference.err; echo "rc=$?"
step "bench c3 / c1 / c4"
python bench.py --workload c3 --steps 3 --warmup 1 > $O/${TAG}_bench_c3.json 2> $O/${TAG}_bench_c3.err; echo "c3 rc=$?"
python bench.py --workload c1 > $O/${TAG}_bench_c1.json 2> $O/${TAG}_bench_c1.err; echo "c1 rc=$?"
python bench.py --workload c4 > $O/${TAG}_bench_c4.json 2> $O/${TAG}_bench_c4.err; echo "c4 rc=$?"
step "ncu k_batch_solve (--set full) + launch list c5"
bash tools/ncu_c5.sh $TAG > $O/${TAG}_ncu_c5_steps.log 2>&1; tail -3 $O/${TAG}_ncu_c5_steps.log
step "ncu c3: launch list + --set full of the frontal and solve kernels (2 loop passes)"
python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 12 --no-cpu-baseline > $O/${TAG}_plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file $O/${TAG}_launches_c3.csv \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 12 --no-cpu-baseline > $O/${TAG}_ncu_c3.log 2>&1; echo "rc=$?"
# launches per Newton step from the launch list (24 Newton steps in that run: 2 solves x 12 passes); the captures below take the
# SECOND Newton step's launches (-s skips the first)
NSOLVE=$(python - <<PY
import csv
rows=list(csv.reader(open("$O/${TAG}_launches_c3.csv")))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]; kn=rows[hi].index('Kernel Name')
n=sum(1 for r in rows[hi+1:] if len(r)>kn and ('k_fwd' in r[kn] or 'k_bwd' in r[kn]))
print(n//24)
PY
)
echo "solve launches per Newton step: $NSOLVE"
ncu --set full --clock-control none -k 'regex:k_fwd|k_bwd' -s $NSOLVE -c $NSOLVE -f -o $O/${TAG}_c3_solve \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 3 --no-cpu-baseline > $O/${TAG}_ncu_c3solve.log 2>&1; echo "rc=$?"
[ -f $O/${TAG}_c3_solve.ncu-rep ] && ncu -i $O/${TAG}_c3_solve.ncu-rep --page raw --csv > $O/${TAG}_c3_solve.csv 2>/dev/null && rm -f $O/${TAG}_c3_solve.ncu-rep
ncu --set full --clock-control none -k 'regex:k_big_panel|k_big_assemble|k_front_smem' -s 112 -c 112 -f -o $O/${TAG}_c3_front \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 3 --no-cpu-baseline > $O/${TAG}_ncu_c3front.log 2>&1; echo "rc=$?"
[ -f $O/${TAG}_c3_front.ncu-rep ] && ncu -i $O/${TAG}_c3_front.ncu-rep --page raw --csv > $O/${TAG}_c3_front.csv 2>/dev/null && rm -f $O/${TAG}_c3_front.ncu-rep
ls -la $O | grep ${TAG}_ | tail -40
