#!/bin/bash
# A/B of intra-level concurrency (forked streams inside the captured graphs)
for c in 0 1; do echo "== QPDO_B200_CONCURRENT=$c"; QPDO_B200_CONCURRENT=$c python tools/factor_check.py c1 cont60 cont100 cont300 2>&1 | grep -E "^c1 |^cont"; done | tee gpurun_out/exp_conc.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/exp_conc.log
