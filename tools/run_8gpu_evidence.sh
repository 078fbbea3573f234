#!/bin/bash
# 8-GPU evidence of one round: the 33-qubit <psi|H|psi> capacity run (2^30 amplitudes per GPU), the default bench at N = 8 and the
# sharded GPU tests.  Run on the GPU box from the repo root:  bash tools/run_8gpu_evidence.sh <tag> [gpus=8] [qubits per GPU of the capacity run=30]
tag=${1:-r3d}
G=${2:-8}
Q=${3:-30}
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29521"
timeout 400 $T bench.py --gpus $G --qubits $Q --no-e2e --no-extra --steps 1 --warmup 3 > gpurun_out/${tag}_expect_33q_${G}gpu.json 2> gpurun_out/${tag}_expect_33q_${G}gpu.err
echo "33q rc=$?"; tail -c 1500 gpurun_out/${tag}_expect_33q_${G}gpu.json; tail -3 gpurun_out/${tag}_expect_33q_${G}gpu.err
timeout 500 $T bench.py --gpus $G > gpurun_out/${tag}_bench_${G}gpu.json 2> gpurun_out/${tag}_bench_${G}gpu.err
echo "bench8 rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_bench_${G}gpu.json"))
print(d["value"], d["ms_per_step"], d["check"], d["roofline"]["nvlink"]["exposed_ms_per_step"], d["roofline"]["nvlink"]["exchange"],
      d["extra"]["apply"]["ms"], d["extra"]["apply"]["nvlink"], d["extra"]["lanczos_step"]["ms"], d["memory"], d["clocks"])
PY
timeout 300 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/${tag}_pytest_sharded_${G}gpu.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest_sharded_${G}gpu.log
