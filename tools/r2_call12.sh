#!/bin/bash
# Round 2, call 12 (1 GPU): host-side trace of the CG solves in the time loop (PYN_KSP_TRACE) + adaptive test.
set -u
out=gpurun_out/r2l
mkdir -p "$out"
timeout 600 python -m pytest tests/test_gpu_timestep.py -q -k adaptive > "$out/pytest.log" 2>&1; echo "pytest rc=$?" | tee "$out/rc.txt"
PYN_KSP_TRACE=1 timeout 600 python bench.py --steps 3 --warmup 2 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree --no-alt \
   > "$out/bench.json" 2> "$out/trace.txt"; echo "bench rc=$?" | tee -a "$out/rc.txt"
PYN_KSP_TRACE=1 timeout 600 python bench.py --steps 3 --warmup 2 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree --no-alt --pc bjacobi \
   > "$out/bench_bj.json" 2> "$out/trace_bj.txt"; echo "bench bj rc=$?" | tee -a "$out/rc.txt"
tail -3 "$out/pytest.log"; grep -c ksp_trace "$out/trace.txt"; tail -16 "$out/trace.txt"; tail -8 "$out/trace_bj.txt"
python - "$out/bench.json" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("ms/step",d["ms_per_step"],"its",d["config"]["ksp_iterations_per_step"])
PY
