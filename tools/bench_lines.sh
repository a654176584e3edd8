#!/bin/bash
# tools/bench_lines.sh [tag] -- the bench lines of both arms and of every workload on one B200 (no ncu): GPU tests, smoke, then the lines
TAG=${1:-r2e}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -1 $O/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > $O/${TAG}_smoke.log 2>&1; tail -1 $O/${TAG}_smoke.log
python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench_c5.json 2> $O/${TAG}_bench_c5.err; echo "c5 rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_c5_reference.json 2> $O/${TAG}_bench_c5_reference.err; echo "ref rc=$?"
python bench.py --workload c3 --steps 3 --warmup 1 > $O/${TAG}_bench_c3.json 2> $O/${TAG}_bench_c3.err; echo "c3 rc=$?"
python bench.py --workload c1 > $O/${TAG}_bench_c1.json 2> $O/${TAG}_bench_c1.err; echo "c1 rc=$?"
python bench.py --workload c4 > $O/${TAG}_bench_c4.json 2> $O/${TAG}_bench_c4.err; echo "c4 rc=$?"
