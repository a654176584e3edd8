#!/bin/bash
# tools/profile_round.sh [tag] -- the round's evidence run on one B200 (under gpurun): GPU tests, smoke, the
# bench lines of both arms and both workloads, then (only after each plain command exited 0) the ncu launch
# lists and the --set full captures.  Everything lands in gpurun_out/<tag>_*.
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
cd "$(dirname "$0")/.."
step() { echo "== $*" | tee -a $O/${TAG}_steps.log; }
: > $O/${TAG}_steps.log

step "pytest -m gpu"
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "rc=$?" | tee -a $O/${TAG}_steps.log
tail -3 $O/${TAG}_pytest_gpu.log
step "smoke"
python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > $O/${TAG}_smoke.log 2>&1; tail -2 $O/${TAG}_smoke.log

step "bench c5 (default)"
python bench.py > $O/${TAG}_bench_c5.json 2> $O/${TAG}_bench_c5.err; echo "rc=$?"
step "bench c5 --impl reference"
python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_c5_reference.json 2> $O/${TAG}_bench_c5_reference.err; echo "rc=$?"
C3RC=1
if [ "$SKIP_C3" != 1 ]; then
step "bench c3"
python bench.py --workload c3 --steps 2 --warmup 1 > $O/${TAG}_bench_c3.json 2> $O/${TAG}_bench_c3.err; C3RC=$?; echo "rc=$C3RC"
fi
if [ "$SKIP_C3_REF" != 1 ] && [ "$SKIP_C3" != 1 ]; then
step "bench c3 --impl reference"
python bench.py --workload c3 --impl reference --steps 1 --warmup 0 > $O/${TAG}_bench_c3_reference.json 2> $O/${TAG}_bench_c3_reference.err; echo "rc=$?"
fi

step "ncu launch list c5"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_plain_c5.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/${TAG}_launches_c5.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_c5.log 2>&1; echo "rc=$?"
step "ncu --set full k_batch_solve"
python tools/batch_time.py 18944 > $O/${TAG}_plain_bt.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_batch_solve -s 8 -c 1 -f -o $O/${TAG}_batch \
    python tools/batch_time.py 18944 > $O/${TAG}_ncu_bt.log 2>&1; echo "rc=$?"

if [ $C3RC = 0 ]; then
step "ncu launch list c3 (8 Newton steps, warm caches)"
python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 12 --no-cpu-baseline > $O/${TAG}_plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file $O/${TAG}_launches_c3.csv \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 12 --no-cpu-baseline > $O/${TAG}_ncu_c3.log 2>&1; echo "rc=$?"
step "ncu --set full: bandwidth kernels of c3 (SpMV, Ruiz, residuals, linesearch build / radix sort / sweep, K assembly)"
ncu --set full --clock-control none \
    -k 'regex:k_spmv_rows|k_spmv_sym|k_ruiz_scaleA|k_residuals|k_assemble_K|k_ls_build|k_sort_hist|k_sort_scan|k_sort_scatter|k_sweep_sums|k_sweep_find|k_update_iterate' \
    -c 110 -f -o $O/${TAG}_c3_bw \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 3 --no-cpu-baseline > $O/${TAG}_ncu_c3bw.log 2>&1; echo "rc=$?"
step "ncu --set full: frontal kernels of c3 (big fronts: assemble + fused panel; small fronts; solves)"
ncu --set full --clock-control none -k 'regex:k_big_panel|k_big_assemble' -c 50 -f -o $O/${TAG}_c3_big \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 2 --no-cpu-baseline > $O/${TAG}_ncu_c3big.log 2>&1; echo "rc=$?"
ncu --set full --clock-control none -k 'regex:k_front_smem|k_fwd|k_bwd' -c 60 -f -o $O/${TAG}_c3_front \
    python bench.py --workload c3 --steps 1 --warmup 0 --max-iter 2 --no-cpu-baseline > $O/${TAG}_ncu_c3front.log 2>&1; echo "rc=$?"
for r in c3_bw c3_big c3_front; do      # keep the raw pages as CSV; the reports themselves are too big to bring back
  [ -f $O/${TAG}_$r.ncu-rep ] && ncu -i $O/${TAG}_$r.ncu-rep --page raw --csv > $O/${TAG}_$r.csv 2>/dev/null && rm -f $O/${TAG}_$r.ncu-rep
done
fi
ls -la $O | tail -30
