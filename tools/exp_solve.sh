#!/bin/bash
# A/B of the cluster solves of the big fronts
for m in cta cluster; do echo "== QPDO_B200_SOLVE=$m"; QPDO_B200_SOLVE=$m python tools/factor_check.py c1 cont100 cont300 2>&1 | grep -E "^c1 |^cont" | sed "s/ n=.*| resid/ resid/"; done | tee gpurun_out/exp_solve.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/exp_solve.log
