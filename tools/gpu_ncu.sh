#!/bin/bash
# ncu passes for the bench command (run only after the same command exited 0 without ncu)
# usage: tools/gpu_ncu.sh [tag]   -> gpurun_out/launches_<tag>.csv (+ prof_audit_<tag>.ncu-rep with FULL=1)
TAG=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs --sessions 10000"
$CMD > gpurun_out/ncu_plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"
if [ -n "$FULL" ]; then
  $CMD > gpurun_out/ncu_plain2_$TAG.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'k_sha256_fixed|k_merkle_level' -s 12 -c 4 -f -o gpurun_out/prof_audit_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
  echo "full rc=$?"
fi
ls -la gpurun_out | head -30
