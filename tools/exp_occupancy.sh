# tools/exp_occupancy.sh -- occupancy experiments for the batch kernel (run on the GPU box)
sh build.sh > /dev/null 2>&1
for pad in 0 20000 60000 150000; do
  QPDO_BATCH_PAD_SMEM=$pad python bench.py --workload c5 --steps 2 --no-cpu-baseline > gpurun_out/exp_pad_$pad.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/exp_pad_$pad.json')); print('pad $pad:', round(d['value']), 'QP/s kernel', round(d['kernel_ms_per_step'],1), 'ms', d['batch_kernel']['smem_bytes'], d['batch_kernel']['ctas'])"
done
