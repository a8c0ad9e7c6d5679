# usage: tools/final_scale.sh [tag]   (one gpurun --gpus 8 session) -> gpurun_out/bench_<tag>_n{8,4,2}.json
TAG=${1:-r02c}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
for N in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+N)) bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/bench_${TAG}_n$N.json 2> gpurun_out/bench_${TAG}_n$N.err; echo "N=$N rc=$?"
done
timeout 300 python -m pytest tests/test_gpu_comm.py -m gpu -x -q 2>&1 | tail -2
python - <<PY
import json
for n in (8,4,2):
    d=json.load(open(f"gpurun_out/bench_${TAG}_n{n}.json")); print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["h2d_ceiling_gbs"], d["e2e"].get("both_directions_over_h2d_ceiling", d["e2e"].get("frac_of_ceiling_both_directions")), d["verified"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
