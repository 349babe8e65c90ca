#!/bin/bash
# Round 2, call 5 (1 GPU): separate the effects of call 4 -- SpMV kernel restored for one GPU, coarse
# grid restored; adaptive chunks on / off.
set -u
out=gpurun_out/r2e
mkdir -p "$out"
run() { tag=$1; shift; timeout 600 python bench.py --steps 5 --warmup 3 --no-asm-warm --no-cpu-baseline --no-e2e "$@" \
          > "$out/bench_$tag.json" 2> "$out/bench_$tag.err"; echo "bench $tag rc=$?" | tee -a "$out/rc.txt"; }
: > "$out/rc.txt"
run bj_adapt --pc bjacobi
run bj_fixed --pc bjacobi --tune ksp_adaptive_chunks=0
run c5_adapt --pc asm --opt pc_asm_coarse=5
run c5_fixed --pc asm --opt pc_asm_coarse=5 --tune ksp_adaptive_chunks=0
run c5_fixed4 --pc asm --opt pc_asm_coarse=5 --tune ksp_adaptive_chunks=0 --check-every 4
B="python bench.py --steps 1 --warmup 1 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv \
   --log-file "$out/launches_c5.csv" $B --pc asm --opt pc_asm_coarse=5 > "$out/ncu_l.log" 2>&1
for f in "$out"/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 2), "its/step", d["config"]["ksp_iterations_per_step"],
          "spmv us", d["roofline"]["avg_launch_us"], "matfree", round(d.get("matfree", {}).get("ms_per_step", 0), 2), "launches", d["gpu_launches"])
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
