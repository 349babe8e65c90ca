#!/bin/bash
set -u
out=gpurun_out/r2n
mkdir -p "$out"
PYN_KSP_TRACE=1 timeout 600 python bench.py --steps 3 --warmup 2 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree --no-alt \
   > "$out/bench_trace.json" 2> "$out/trace.txt"; echo "bench trace rc=$?" | tee "$out/rc.txt"
timeout 600 python bench.py --steps 10 --warmup 3 --no-asm-warm --no-cpu-baseline > "$out/bench.json" 2> "$out/bench.err"; echo "bench rc=$?" | tee -a "$out/rc.txt"
timeout 900 python -m pytest tests/test_gpu_timestep.py tests/test_gpu_solve.py tests/test_gpu_asm.py -q -x > "$out/pytest.log" 2>&1; echo "pytest rc=$?" | tee -a "$out/rc.txt"
tail -9 "$out/trace.txt"; tail -3 "$out/pytest.log"
python - "$out/bench.json" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("ms/step",round(d["ms_per_step"],2),"e2e",round(d["e2e"]["ms_per_step"],2),"its",d["config"]["ksp_iterations_per_step"],"matfree",round(d["matfree"]["ms_per_step"],2),"alt",round(d["alt_pc_bjacobi"]["ms_per_step"],2),"launches",d["gpu_launches"])
PY
