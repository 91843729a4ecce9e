#!/bin/bash
# One GPU-box visit: bench line, GPU test suite, A/B timings of the cache-hint builds.  Outputs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > gpurun_out/gpu.txt 2>&1
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench.json
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/ab.log
for v in base h7 h1 h2 h4 base; do
  if [ $v = base ]; then L=total-lagrangian-fea_b200/lib/libtlfea_b200.so; else L=total-lagrangian-fea_b200/lib/ab/libtlfea_$v.so; fi
  echo "== $v" >> gpurun_out/ab.log
  TLFEA_LIB=$PWD/$L timeout 120 python tools/time_assemble.py 119 1 >> gpurun_out/ab.log 2>&1
done
cat gpurun_out/ab.log
