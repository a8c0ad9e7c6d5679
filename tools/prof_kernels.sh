#!/bin/bash
# ncu --set full of the kernels VERDICT r01 found unprofiled (run only after the plain run exited 0).
# usage: tools/prof_kernels.sh <round-tag> [sigma|var|chain|tree]   -> gpurun_out/prof_<tag>_<kernel>.ncu-rep
TAG=${1:-r02}
mkdir -p gpurun_out
python tools/prof_kernels.py > gpurun_out/prof_plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain_$TAG.log; exit 1; }
ONLY=${2:-all}
prof() {   # name, workload, kernel regex, skip, count
  [ "$ONLY" = all ] || [ "$ONLY" = "$1" ] || return 0
  ncu --set full --clock-control none --import-source on -k regex:"$3" -s $4 -c $5 -f -o gpurun_out/prof_${TAG}_$1 \
      python tools/prof_kernels.py $2 > gpurun_out/prof_${TAG}_$1.log 2>&1
  echo "$1 rc=$?"
}
prof sigma sigma 'k_sigma_eff_tiled' 2 2          # launch 3 = Zipf, launch 4 = uniform
prof var var 'k_sha256_var' 1 1
prof chain chain 'k_chain_prefix|k_chain_capture' 2 2      # second call: the pre-pass and the chain walk
prof tree tree 'k_merkle_level|k_merkle_top' 20 2   # levels 1 and 2 of a uniform 10^4 x 1000 tree
# the slashing lane kernel on the mid-size batch (tools/slash_lab.py picks the host entry's slice)
python tools/slash_lab.py --sessions 500000 --only 32,256 > gpurun_out/prof_plain_slash_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_slash_lanes -s 2 -c 1 -f -o gpurun_out/prof_${TAG}_slash_mid \
    python tools/slash_lab.py --sessions 500000 --only 32,256 > gpurun_out/prof_${TAG}_slash_mid.log 2>&1
echo "slash_mid rc=$?"
du -sh gpurun_out; ls -la gpurun_out | head -40   # gpurun copies back at most 64 MiB
