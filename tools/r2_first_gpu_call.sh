#!/bin/bash
# First GPU call of the next round: confirm -pc_type asm (written and CPU-emulated in round 1 after the
# GPU minutes ran out) and measure it on the bench workload.  Run as
#   gpurun --timeout 1500 -- 'bash tools/r2_first_gpu_call.sh'
# Everything lands in gpurun_out/r2_asm/.  Then, on 2 GPUs (gpurun --gpus 2): the same pytest line runs
# test_partitioned_run_with_asm, and
#   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 \
#       bench.py --gpus 2 --steps 3 --warmup 3 --pc asm
set -u
out=gpurun_out/r2_asm
mkdir -p "$out"
# memory errors first, on the smallest case (these kernels have never touched a GPU)
PYN_TEST_ASM=1 timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 \
  python -m pytest tests/test_gpu_asm.py -x -q -k "reproducible" > "$out/memcheck.log" 2>&1
echo "memcheck rc=$?" | tee "$out/rc.txt"
PYN_TEST_ASM=1 timeout 900 python -m pytest tests/test_gpu_asm.py -x -q > "$out/pytest_asm.log" 2>&1
echo "pytest_asm rc=$?" | tee -a "$out/rc.txt"
for cfg in "bjacobi" "asm" "asm --opt pc_asm_coarse=5"; do
  tag=$(echo "$cfg" | tr -c 'a-z0-9' '_')
  timeout 600 python bench.py --steps 3 --warmup 3 --pc $cfg --no-asm-warm > "$out/bench_$tag.json" 2> "$out/bench_$tag.err"
  echo "bench $cfg rc=$?" | tee -a "$out/rc.txt"
done
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
