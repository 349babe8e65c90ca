#!/bin/bash
# Round 2, call 11 (8 GPUs): multi-rank parity tests at 4 and 8 ranks, strong scaling of the bench at 4 and 8
# GPUs (default = two-level Schwarz; node-block Jacobi beside it), configs[3] (cylinder) on 8 GPUs,
# configs[4] sweep points on 8 GPUs.  Charged 8x: keep it short.
set -u
out=gpurun_out/r2k
mkdir -p "$out"
nvidia-smi -L | wc -l > "$out/ngpus.txt"
T8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
T4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
: > "$out/rc.txt"
timeout 600 python -m pytest tests/test_gpu_dist.py tests/test_gpu_asm.py -q \
   -k "matches_oracle or cylinder_channel_partitioned or two_level_asm or (cavity and 4)" > "$out/pytest_dist.log" 2>&1
echo "pytest_dist rc=$?" | tee -a "$out/rc.txt"
B="bench.py --steps 10 --warmup 3 --no-asm-warm --no-e2e"
timeout 300 $T8 --master-port 29621 $B --gpus 8 > "$out/bench_n8.json" 2> "$out/bench_n8.err"; echo "bench n8 rc=$?" | tee -a "$out/rc.txt"
timeout 300 $T4 --master-port 29622 $B --gpus 4 > "$out/bench_n4.json" 2> "$out/bench_n4.err"; echo "bench n4 rc=$?" | tee -a "$out/rc.txt"
timeout 300 $T8 --master-port 29623 $B --gpus 8 --pc bjacobi --no-matfree > "$out/bench_n8_bjacobi.json" 2> "$out/bench_n8_bjacobi.err"; echo "bench n8 bjacobi rc=$?" | tee -a "$out/rc.txt"
timeout 300 $T8 --master-port 29624 $B --gpus 8 --pc bjacobi --no-matfree --tune halo_wait_at_start=0 > "$out/bench_n8_bjacobi_deferred.json" 2> "$out/bench_n8_bjacobi_deferred.err"; echo "bench n8 bjacobi deferred rc=$?" | tee -a "$out/rc.txt"
timeout 300 $T8 --master-port 29625 bench.py --steps 5 --warmup 2 --no-asm-warm --no-e2e --nelem 32 --gpus 8 > "$out/bench_32cube_n8.json" 2> "$out/bench_32cube_n8.err"; echo "bench 32^3 n8 rc=$?" | tee -a "$out/rc.txt"
timeout 600 $T8 --master-port 29626 tools/run_cylinder.py --h 0.25 --steps 2 > "$out/cylinder_n8.json" 2> "$out/cylinder_n8.err"; echo "cylinder n8 rc=$?" | tee -a "$out/rc.txt"
timeout 900 $T8 --master-port 29627 tools/sweep_dist.py "32,32,32,4 64,64,64,3 128,128,128,3 64,64,64,4 128,128,128,4 64,64,64,5" > "$out/sweep_n8.jsonl" 2> "$out/sweep_n8.err"; echo "sweep n8 rc=$?" | tee -a "$out/rc.txt"
tail -6 "$out/pytest_dist.log"
for f in "$out"/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 2), "its/step", d["config"]["ksp_iterations_per_step"],
          "spmv us", d["roofline"]["avg_launch_us"], "frac", d["roofline"]["frac"], "matfree", round((d.get("matfree") or {}).get("ms_per_step", 0), 2), d["checksum"]["vort_l2"])
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
cut -c1-700 "$out/cylinder_n8.json" "$out/sweep_n8.jsonl"
tail -3 "$out/cylinder_n8.err" "$out/sweep_n8.err"
cat "$out/rc.txt"
