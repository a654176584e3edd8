#!/bin/bash
# tools/ncu_c5.sh [tag] -- on the GPU box: launch list of the default bench + one `ncu --set full` capture of a
# k_batch_solve launch of 2,368 QPs (4 per resident CTA); raw page as CSV in gpurun_out/<tag>_batch.csv
TAG=${1:-r2}
O=gpurun_out
cd "$(dirname "$0")/.."
python tools/batch_time.py 2368 > $O/${TAG}_plain_bt.log 2>&1 || { tail -3 $O/${TAG}_plain_bt.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:k_batch_solve -s 2 -c 1 -f -o $O/${TAG}_batch \
    python tools/batch_time.py 2368 > $O/${TAG}_ncu_bt.log 2>&1; echo "ncu rc=$?"
ncu -i $O/${TAG}_batch.ncu-rep --page raw --csv > $O/${TAG}_batch.csv 2>/dev/null
python tools/ncu_lines.py $O/${TAG}_batch.ncu-rep > $O/${TAG}_batch_lines.txt 2>/dev/null || true
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kkt > $O/${TAG}_plain_c5.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/${TAG}_launches_c5.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kkt > $O/${TAG}_ncu_c5.log 2>&1; echo "launch list rc=$?"
ls -la $O | grep ${TAG}_
