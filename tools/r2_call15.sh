#!/bin/bash
set -u
out=gpurun_out/r2p
mkdir -p "$out"
timeout 900 python -m pytest tests/test_gpu_assembly.py tests/test_gpu_fullsize.py tests/test_gpu_solve.py -q -x > "$out/pytest.log" 2>&1; echo "pytest rc=$?" | tee "$out/rc.txt"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-alt --no-matfree --no-e2e > "$out/bench.json" 2> "$out/bench.err"; echo "bench rc=$?" | tee -a "$out/rc.txt"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-alt --no-matfree --no-e2e --tune assemble_row_classes=0 > "$out/bench_off.json" 2> "$out/bench_off.err"; echo "bench off rc=$?" | tee -a "$out/rc.txt"
tail -3 "$out/pytest.log"
for f in "$out/bench.json" "$out/bench_off.json"; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split("/")[-1], json.dumps(d["assembly_kernels_warm"])[:900]); print("  first assembly", d["assembly"]["phase_ms"], d["assembly"]["kernels"])
PY
done
