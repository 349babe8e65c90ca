#!/bin/bash
# Round 2, call 8 (2 GPUs): two-level Schwarz preconditioner on a row-partitioned operator (replicated
# coarse problem) -- tests at 1 and 2 ranks, benches at N=1 and N=2, launch list of the time loop.
set -u
out=gpurun_out/r2h
mkdir -p "$out"
timeout 900 python -m pytest tests/test_gpu_asm.py -q > "$out/pytest_asm.log" 2>&1
echo "pytest_asm rc=$?" | tee "$out/rc.txt"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
B="bench.py --steps 10 --warmup 3 --no-asm-warm --no-matfree --no-e2e --no-cpu-baseline --pc asm --opt pc_asm_coarse=5"
timeout 600 $T --master-port 29611 $B --gpus 2 > "$out/bench_n2_c5.json" 2> "$out/bench_n2_c5.err"
echo "bench n2 c5 rc=$?" | tee -a "$out/rc.txt"
timeout 600 python $B > "$out/bench_n1_c5.json" 2> "$out/bench_n1_c5.err"
echo "bench n1 c5 rc=$?" | tee -a "$out/rc.txt"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 2600 -c 2500 --csv \
   --log-file "$out/launches_loop_c5.csv" python bench.py --steps 1 --warmup 1 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree \
   --pc asm --opt pc_asm_coarse=5 > "$out/ncu_l.log" 2>&1
echo "launch list rc=$?" | tee -a "$out/rc.txt"
tail -6 "$out/pytest_asm.log"
for f in "$out"/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 2), "its/step", d["config"]["ksp_iterations_per_step"],
          "spmv us", d["roofline"]["avg_launch_us"], "frac", d["roofline"]["frac"], d.get("checksum"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
tail -5 "$out"/bench_n2_c5.err
cat "$out/rc.txt"
