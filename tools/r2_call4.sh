#!/bin/bash
# Round 2, call 4 (1 GPU): whole GPU suite after the SpMV row-order / chunking / coarse-grid changes, then
# N=1 benches: bjacobi (SpMV regression check), asm, asm + coarse.
set -u
out=gpurun_out/r2d
mkdir -p "$out"
timeout 1500 python -m pytest tests -m gpu -x -q > "$out/pytest_gpu.log" 2>&1
echo "pytest_gpu rc=$?" | tee "$out/rc.txt"
for cfg in "bjacobi" "asm" "asm --opt pc_asm_coarse=5" "asm --opt pc_asm_coarse=4"; do
  tag=$(echo "$cfg" | tr -c 'a-z0-9' '_')
  timeout 600 python bench.py --steps 5 --warmup 3 --pc $cfg --no-asm-warm --no-cpu-baseline \
      > "$out/bench_$tag.json" 2> "$out/bench_$tag.err"
  echo "bench $cfg rc=$?" | tee -a "$out/rc.txt"
done
tail -4 "$out/pytest_gpu.log"
for f in "$out"/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 2), "its/step", d["config"]["ksp_iterations_per_step"],
          "spmv us", d["roofline"]["avg_launch_us"], "frac", d["roofline"]["frac"], "matfree", round(d.get("matfree", {}).get("ms_per_step", 0), 2),
          "launches", d["gpu_launches"])
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
cat "$out/rc.txt"
