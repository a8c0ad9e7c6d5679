#!/bin/bash
# One multi-GPU session (gpurun --gpus N): copy ceiling probe, bench at N and N/2, the 10^8-leaf tree, comm test.
# usage: tools/multi_gpu_run.sh N tag
N=${1:-8}; TAG=${2:-r02}
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
timeout 300 python -m pytest tests/test_gpu_comm.py -m gpu -x -q 2>&1 | tail -3
timeout 600 bash -c "$(declare -f run); run $N 29601 tools/h2d_probe.py" > gpurun_out/h2d_probe_n${N}_$TAG.out 2> gpurun_out/h2d_probe_n${N}_$TAG.err
tail -n 1 gpurun_out/h2d_probe_n${N}_$TAG.out > gpurun_out/h2d_probe_n${N}_$TAG.json; echo "probe rc=$?"
timeout 400 bash -c "$(declare -f run); run $N 29602 bench.py --gpus $N --steps 20 --warmup 3" > gpurun_out/bench_${TAG}_n$N.json 2> gpurun_out/bench_${TAG}_n$N.err; echo "bench N=$N rc=$?"
H=$((N/2))
if [ $H -ge 2 ]; then
  timeout 400 bash -c "$(declare -f run); run $H 29603 bench.py --gpus $H --steps 20 --warmup 3" > gpurun_out/bench_${TAG}_n$H.json 2> gpurun_out/bench_${TAG}_n$H.err; echo "bench N=$H rc=$?"
fi
timeout 600 bash -c "$(declare -f run); run $N 29604 tools/giant_tree.py" > gpurun_out/giant_tree_n${N}_$TAG.out 2> gpurun_out/giant_tree_n${N}_$TAG.err; echo "giant rc=$?"
tail -n 1 gpurun_out/giant_tree_n${N}_$TAG.out
python - <<PY
import json
for n in ($N, $H):
    try:
        d = json.load(open(f"gpurun_out/bench_${TAG}_n{n}.json"))
        print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["h2d_ceiling_gbs"], d["e2e"]["frac_of_h2d_ceiling"], d["verified"], d["comm"])
    except Exception as e:
        print(n, "no bench line", e)
try:
    d = json.load(open("gpurun_out/h2d_probe_n${N}_$TAG.json"))
    for k, row in d["by_active_ranks"].items():
        print(k, {m: round(v["GBps_aggregate"], 1) for m, v in row.items()})
except Exception as e:
    print("no probe", e)
PY
