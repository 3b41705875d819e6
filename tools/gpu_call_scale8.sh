#!/bin/bash
# 8-GPU call: probe with the data-path sweep, then bench.py in the optimizer placements / data paths to choose from.
N=${1:-8}; TAG=${2:-s8}
export B200_P2P_TIMEOUT_S=${B200_P2P_TIMEOUT_S:-30} B200_COMM_TIMEOUT_S=60
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout -k 10 400 $TR --master-port 29511 tools/gpu_probe_p2p.py > gpurun_out/probe_${N}gpu_${TAG}.log 2>&1; echo "probe rc=$?"
grep -E "PROBE|BW|SWEEP|sharded|allreduce step|comm runtime" gpurun_out/probe_${N}gpu_${TAG}.log | tail -150
run() {  # name, env...
  name=$1; shift
  env "$@" timeout -k 10 300 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu_${name}_${TAG}.log 2>&1; echo "bench $name rc=$?"
  tail -1 gpurun_out/bench_${N}gpu_${name}_${TAG}.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print({k:d[k] for k in ('value','ms_per_step','final_loss','exposed_allreduce_ms_per_step','replicas_identical') if k in d}, d.get('e2e',{}).get('value'))" || tail -5 gpurun_out/bench_${N}gpu_${name}_${TAG}.log
}
run pipelined B200_PIPELINED_UPDATE=1
run nodefer X=1
run nodefer_peer B200_RS_PEER=1 B200_BCAST_PEER=1
run nodefer_bcastpeer B200_BCAST_PEER=1
run pipelined_bcastpeer B200_PIPELINED_UPDATE=1 B200_BCAST_PEER=1
run replicated B200_NO_SHARDED=1
