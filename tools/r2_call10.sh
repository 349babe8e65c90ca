#!/bin/bash
# Round 2, call 10 (2 GPUs): peer-memory reductions under the Schwarz preconditioner; adaptive-step test.
set -u
out=gpurun_out/r2j
mkdir -p "$out"
timeout 900 python -m pytest tests/test_gpu_asm.py tests/test_gpu_timestep.py -q -k "partitioned or adaptive" > "$out/pytest.log" 2>&1
echo "pytest rc=$?" | tee "$out/rc.txt"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $T --master-port 29611 bench.py --gpus 2 --steps 10 --warmup 3 --no-asm-warm > "$out/bench_n2.json" 2> "$out/bench_n2.err"
echo "bench n2 rc=$?" | tee -a "$out/rc.txt"
PYN_P2P=0 timeout 600 $T --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 --no-asm-warm --no-matfree --no-e2e > "$out/bench_n2_nccl.json" 2> "$out/bench_n2_nccl.err"
echo "bench n2 nccl rc=$?" | tee -a "$out/rc.txt"
tail -5 "$out/pytest.log"
for f in "$out"/bench_n*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 2), "its/step", d["config"]["ksp_iterations_per_step"],
          "spmv us", d["roofline"]["avg_launch_us"], "matfree", round((d.get("matfree") or {}).get("ms_per_step", 0), 2), d["checksum"], d["config"]["parallelism"][-60:])
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
tail -4 "$out/bench_n2.err"; cat "$out/rc.txt"
