#!/bin/bash
# Round 2, call 7 (2 GPUs): owned nodes renumbered interior-first, halo wait deferred without indirection.
set -u
out=gpurun_out/r2g
mkdir -p "$out"
timeout 900 python -m pytest tests/test_gpu_dist.py -q -x -k "partitioned_run_matches or single_reduction or cylinder" > "$out/pytest_dist.log" 2>&1
echo "pytest_dist rc=$?" | tee "$out/rc.txt"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
B="bench.py --gpus 2 --steps 10 --warmup 3 --no-asm-warm --no-matfree --no-e2e"
timeout 600 $T --master-port 29611 $B > "$out/bench_n2_deferred.json" 2> "$out/bench_n2_deferred.err"
echo "bench n2 deferred rc=$?" | tee -a "$out/rc.txt"
timeout 600 $T --master-port 29612 $B --tune halo_wait_at_start=1 > "$out/bench_n2_start.json" 2> "$out/bench_n2_start.err"
echo "bench n2 at-start rc=$?" | tee -a "$out/rc.txt"
tail -4 "$out/pytest_dist.log"
for f in "$out"/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 2), "its/step", d["config"]["ksp_iterations_per_step"],
          "spmv us", d["roofline"]["avg_launch_us"], "frac", d["roofline"]["frac"], d["roofline"].get("over_ranks"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
cat "$out/rc.txt"
