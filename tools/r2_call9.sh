#!/bin/bash
# Round 2, call 9 (2 GPUs): dry run of what the 8-GPU call will do -- bench default (asm + coarse) at N=1
# (with the bjacobi comparison block and the CPU baseline) and N=2; the sweep tool at 1 and 2 ranks.
set -u
out=gpurun_out/r2i
mkdir -p "$out"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python bench.py --steps 10 --warmup 3 > "$out/bench_n1.json" 2> "$out/bench_n1.err"
echo "bench n1 rc=$?" | tee "$out/rc.txt"
timeout 600 $T --master-port 29611 bench.py --gpus 2 --steps 10 --warmup 3 > "$out/bench_n2.json" 2> "$out/bench_n2.err"
echo "bench n2 rc=$?" | tee -a "$out/rc.txt"
timeout 600 python tools/sweep_dist.py "16,16,16,4 32,32,32,3 32,32,32,4 64,64,64,6" > "$out/sweep_n1.jsonl" 2> "$out/sweep_n1.err"
echo "sweep n1 rc=$?" | tee -a "$out/rc.txt"
timeout 600 $T --master-port 29612 tools/sweep_dist.py "16,16,16,4 32,32,32,3 64,32,32,3 32,32,32,4" > "$out/sweep_n2.jsonl" 2> "$out/sweep_n2.err"
echo "sweep n2 rc=$?" | tee -a "$out/rc.txt"
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > "$out/bench_ref.json" 2> "$out/bench_ref.err"
echo "bench ref rc=$?" | tee -a "$out/rc.txt"
nproc > "$out/nproc.txt"; free -g | head -2 >> "$out/nproc.txt"
for f in "$out"/bench_n*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["ms_per_step"], 2), "its/step", d["config"]["ksp_iterations_per_step"],
          "spmv us", d["roofline"]["avg_launch_us"], "frac", d["roofline"]["frac"], "matfree", round(d.get("matfree", {}).get("ms_per_step", 0), 2),
          "alt", d.get("alt_pc_bjacobi", {}).get("ms_per_step"), "fp64", d.get("fp64_peak", {}).get("TFLOPs"))
    print("   cpu", d.get("cpu_baseline"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
cut -c1-600 "$out"/sweep_n1.jsonl "$out"/sweep_n2.jsonl; cut -c1-900 "$out/bench_ref.json"; tail -3 "$out"/*.err | tail -30
cat "$out/rc.txt" "$out/nproc.txt"
