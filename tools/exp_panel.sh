#!/bin/bash
# tools/exp_panel.sh -- A/B of the panel step on one B200: factor_check (residual of K x = b, factor / solve ms) with the
# round-1 panel step (QPDO_B200_PANEL=v1) and the current one, then the GPU tests
O=gpurun_out
mkdir -p $O
for v in v1 v2; do
  echo "== panel $v"
  QPDO_B200_PANEL=$v timeout 600 python tools/factor_check.py c1 cont60 cont100 cont300 2>&1 | tail -8
done | tee $O/exp_panel.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee -a $O/exp_panel.log
