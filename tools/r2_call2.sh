#!/bin/bash
# Round 2, second GPU call: all -pc_type asm tests (no -x: see every failure), then launch lists of the
# bench under asm and asm + coarse level, then full ncu captures of the two preconditioner kernels.
set -u
out=gpurun_out/r2b
mkdir -p "$out"
PYN_TEST_ASM=1 timeout 1200 python -m pytest tests/test_gpu_asm.py -q > "$out/pytest_asm.log" 2>&1
echo "pytest_asm rc=$?" | tee "$out/rc.txt"
B="python bench.py --steps 1 --warmup 1 --no-asm-warm --no-cpu-baseline --no-e2e --no-matfree"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
   --log-file "$out/launches_asm_coarse.csv" $B --pc asm --opt pc_asm_coarse=5 > "$out/ncu_l.log" 2>&1
echo "launch list rc=$?" | tee -a "$out/rc.txt"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:asm_mult_kernel -s 20 -c 1 \
   -o "$out/asm_mult" -f $B --pc asm > "$out/ncu_mult.log" 2>&1
echo "ncu mult rc=$?" | tee -a "$out/rc.txt"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:asm_sum_kernel -s 20 -c 1 \
   -o "$out/asm_sum" -f $B --pc asm > "$out/ncu_sum.log" 2>&1
echo "ncu sum rc=$?" | tee -a "$out/rc.txt"
tail -15 "$out/pytest_asm.log"
cat "$out/rc.txt"; ls -la "$out"
