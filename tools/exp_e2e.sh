for sz in 65536 8192; do for ho in 1 0; do for ch in 0 2 4 8 16; do
QPDO_BATCH_HOST_ORDER=$ho QPDO_BATCH_CHUNKS=$ch python bench.py --size $sz --steps 5 --warmup 3 --no-kkt --no-cpu-baseline > gpurun_out/x.json 2> gpurun_out/x.err; python -c "
import json
d=json.load(open('gpurun_out/x.json'))
print('size $sz host_order $ho chunks $ch:', round(d['value']), 'e2e', round(d['e2e']['value']), 'kernel', round(d['kernel_ms_per_step'],2))
"; done; done; done
