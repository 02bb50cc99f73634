#!/bin/bash
# One multi-GPU lease, every number BASELINE.md quotes for that world size:
#   bash scripts/final_multi.sh N      (under gpurun --gpus N)
# scaling bench (2 fresh launches), the torch arm, the reference arm, config 3 (FedAvg of the full state_dict every 5 local
# steps), and the aggregation sweep on the default heap and on the CUDA-IPC heap. Everything lands in gpurun_out/.
N=${1:-8}
run() { port=$((29600 + RANDOM % 300)); timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
bash scripts/scale_check.sh $N 2
grep '"metric"' gpurun_out/sc.out | tail -1 > gpurun_out/r2_bench_ours_$N.json
run bench.py --gpus $N --impl reference --steps 10 --warmup 3 2>/dev/null | grep '"metric"' | tail -1 > gpurun_out/r2_bench_ref_$N.json
run bench.py --gpus $N --impl torch --steps 30 --warmup 3 2>/dev/null | grep '"metric"' | tail -1 > gpurun_out/r2_bench_torch_$N.json
run bench.py --gpus $N --steps 50 --warmup 5 --mode param_avg --local_steps 5 --average_frozen 2>/dev/null | grep '"metric"' | tail -1 > gpurun_out/r2_bench_ours_${N}_paramavg.json
run bench.py --gpus $N --steps 50 --warmup 5 --transport collective 2>/dev/null | grep '"metric"' | tail -1 > gpurun_out/r2_bench_ours_${N}_nccl.json
run scripts/bench_aggregate.py > gpurun_out/agg_sweep_${N}.log 2>&1
FEDREC_SYMM=ipc run scripts/bench_aggregate.py > gpurun_out/agg_sweep_${N}_ipc.log 2>&1
for f in gpurun_out/r2_bench_*_$N*.json; do echo "$f: $(cut -c1-200 $f)"; done
grep -h '"mb": 4.66\|"mb": 270.1' gpurun_out/agg_sweep_${N}.log gpurun_out/agg_sweep_${N}_ipc.log | cut -c1-420
