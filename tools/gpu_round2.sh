#!/bin/bash
# One GPU-box visit of round 2: GPU tests, bench line, ncu launch list + DRAM traffic of the bench command.
mkdir -p gpurun_out
TAG=${1:-r2}
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > gpurun_out/${TAG}_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/${TAG}_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/${TAG}_bench.json
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
SHORT="python bench.py --steps 2 --warmup 3 --no-ancf --no-newton --no-variants --no-cpu-baseline --no-e2e"
timeout 600 $SHORT > gpurun_out/${TAG}_short.json 2> gpurun_out/${TAG}_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv \
  --log-file gpurun_out/${TAG}_launches.csv $SHORT > gpurun_out/${TAG}_ncu_bench.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/${TAG}_launches.csv
