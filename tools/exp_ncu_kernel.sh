#!/bin/bash
# tools/exp_ncu_kernel.sh <kernel regex> <problem> [skip] [count] -- ncu --set full with source on launches of one kernel in factor_check
K=$1; P=${2:-cont300}; S=${3:-0}; C=${4:-4}
O=gpurun_out
python tools/factor_check.py $P > $O/ncu_k_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c $C -f -o $O/kern python tools/factor_check.py $P > $O/ncu_k.log 2>&1
echo "ncu rc=$?"
python tools/ncu_lines.py $O/kern.ncu-rep 28 > $O/kern_lines.txt 2>&1
ncu -i $O/kern.ncu-rep --page raw --csv > $O/kern_raw.csv 2>/dev/null
cut -c1-230 $O/kern_lines.txt
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/kern_raw.csv")) if len(r)>10]
h=rows[0]; 
def col(n): return h.index(n) if n in h else None
names=["Kernel Name","gpu__time_duration.sum","launch__grid_size","launch__waves_per_multiprocessor","sm__warps_active.avg.pct_of_peak_sustained_active","dram__bytes_read.sum","dram__bytes_write.sum","lts__t_bytes.sum","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active","launch__registers_per_thread","launch__occupancy_limit_shared_mem","launch__occupancy_limit_registers"]
for r in rows[2:]:
    print(" | ".join("%s=%s"%(n.split("__")[-1][:28], r[col(n)][:40]) for n in names if col(n) is not None))
PY
