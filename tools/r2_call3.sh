#!/bin/bash
# Round 2, third GPU call: fused coarse level of -pc_type asm -- tests, bench, launch list; the drop-in
# test of the reference's unmodified callers needs /root/reference, which the box does not have (skips).
set -u
out=gpurun_out/r2c
mkdir -p "$out"
timeout 1200 python -m pytest tests/test_gpu_asm.py tests/test_reference_dropin.py -q > "$out/pytest_asm.log" 2>&1
echo "pytest_asm rc=$?" | tee "$out/rc.txt"
for cfg in "asm --opt pc_asm_coarse=5" "asm --opt pc_asm_coarse=3" "asm --opt pc_asm_coarse=8"; do
  tag=$(echo "$cfg" | tr -c 'a-z0-9' '_')
  timeout 600 python bench.py --steps 5 --warmup 3 --pc $cfg --no-asm-warm --no-cpu-baseline \
      > "$out/bench_$tag.json" 2> "$out/bench_$tag.err"
  echo "bench $cfg rc=$?" | tee -a "$out/rc.txt"
done
B="python bench.py --steps 1 --warmup 1 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
   --log-file "$out/launches_asm_coarse.csv" $B --pc asm --opt pc_asm_coarse=5 > "$out/ncu_l.log" 2>&1
echo "launch list rc=$?" | tee -a "$out/rc.txt"
tail -5 "$out/pytest_asm.log"
for f in "$out"/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["config"]["ksp_options"], "ms/step", round(d["ms_per_step"], 2), "its/step",
          d["config"]["ksp_iterations_per_step"], "matfree", d.get("matfree", {}).get("ms_per_step"), d.get("checksum"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
cat "$out/rc.txt"
