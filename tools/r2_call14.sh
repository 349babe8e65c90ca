#!/bin/bash
set -u
out=gpurun_out/r2o
mkdir -p "$out"
timeout 900 python -m pytest tests/test_gpu_solve.py tests/test_gpu_asm.py tests/test_gpu_timestep.py tests/test_gpu_fullsize.py -q -x > "$out/pytest.log" 2>&1; echo "pytest rc=$?" | tee "$out/rc.txt"
PYN_KSP_TRACE=1 timeout 600 python bench.py --steps 3 --warmup 2 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree --no-alt \
   > "$out/bench_trace.json" 2> "$out/trace.txt"; echo "bench trace rc=$?" | tee -a "$out/rc.txt"
timeout 600 python bench.py --steps 10 --warmup 3 --no-asm-warm --no-cpu-baseline > "$out/bench.json" 2> "$out/bench.err"; echo "bench rc=$?" | tee -a "$out/rc.txt"
timeout 600 python bench.py --steps 10 --warmup 3 --no-asm-warm --no-cpu-baseline --no-alt --no-matfree --no-e2e --tune ksp_graphs=0 > "$out/bench_nograph.json" 2> "$out/bench_nograph.err"; echo "bench nograph rc=$?" | tee -a "$out/rc.txt"
tail -8 "$out/trace.txt"; tail -3 "$out/pytest.log"
for f in "$out/bench.json" "$out/bench_nograph.json"; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split("/")[-1],"ms/step",round(d["ms_per_step"],2),"its",d["config"]["ksp_iterations_per_step"],"matfree",round((d.get("matfree") or {}).get("ms_per_step",0),2),"alt",round((d.get("alt_pc_bjacobi") or {}).get("ms_per_step",0),2),"launches",d["gpu_launches"],d["checksum"]["vort_l2"])
PY
done
