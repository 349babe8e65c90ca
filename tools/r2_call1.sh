#!/bin/bash
# Round 2, first GPU call: the new parity tests, the allocation-free entry point, -pc_type asm for the
# first time on a GPU (memcheck of the smallest case first), and a short bench.
set -u
out=gpurun_out/r2a
mkdir -p "$out"
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > "$out/gpu.txt" 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > "$out/pytest_noalloc_alone.log" 2>&1
echo "noalloc_alone rc=$?" | tee "$out/rc.txt"
timeout 1500 python -m pytest tests -m gpu -x -q -s --durations=15 > "$out/pytest_gpu.log" 2>&1
echo "pytest_gpu rc=$?" | tee -a "$out/rc.txt"
PYN_TEST_ASM=1 timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 \
  python -m pytest tests/test_gpu_asm.py -x -q -k "reproducible" > "$out/memcheck.log" 2>&1
echo "memcheck rc=$?" | tee -a "$out/rc.txt"
PYN_TEST_ASM=1 timeout 900 python -m pytest tests/test_gpu_asm.py -x -q > "$out/pytest_asm.log" 2>&1
echo "pytest_asm rc=$?" | tee -a "$out/rc.txt"
for cfg in "bjacobi" "asm" "asm --opt pc_asm_coarse=5"; do
  tag=$(echo "$cfg" | tr -c 'a-z0-9' '_')
  timeout 600 python bench.py --steps 3 --warmup 3 --pc $cfg --no-asm-warm --no-cpu-baseline \
      > "$out/bench_$tag.json" 2> "$out/bench_$tag.err"
  echo "bench $cfg rc=$?" | tee -a "$out/rc.txt"
done
tail -5 "$out/pytest_gpu.log"
tail -3 "$out/pytest_asm.log"
for f in "$out"/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["config"]["ksp_options"], "ms/step", round(d["ms_per_step"], 2), "its/step",
          d["config"]["ksp_iterations_per_step"], "matfree", d.get("matfree", {}).get("ms_per_step"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
cat "$out/rc.txt"
