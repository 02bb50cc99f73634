#!/bin/bash
# Multi-GPU health check for one world size: the driver's scaling bench command, several fresh launches, every heap
# backend. Usage (under gpurun --gpus N):  bash scripts/scale_check.sh N [launches]
# Writes gpurun_out/scale_check_N.txt (one JSON line or FAIL per launch).
N=${1:-4}; REP=${2:-3}
OUT=gpurun_out/scale_check_${N}.txt
mkdir -p gpurun_out; : > $OUT
run() {   # label, env...
  label=$1; shift
  port=$((29600 + RANDOM % 300))
  env "$@" timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $port bench.py --gpus $N --steps 20 --warmup 5 $EXTRA > gpurun_out/sc.out 2> gpurun_out/sc.err
  rc=$?
  line=$(grep '"metric"' gpurun_out/sc.out | tail -1 | python -c 'import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d["config"]; print(json.dumps({"value":d["value"],"ms":d["ms_per_step"],"e2e":d["e2e"]["value"],"heap":c.get("peer_heap"),"transport":c.get("transport"),"opt":c.get("optimizer","")[:9],"degraded":c.get("degraded")}))')
  echo "$label rc=$rc $line" | tee -a $OUT
  if [ $rc -ne 0 ]; then grep -v "^\s*$" gpurun_out/sc.err | grep -iv "warn\|OMP_NUM\|\*\*\*" | tail -15 | tee -a $OUT; fi
}
for i in $(seq 1 $REP); do run "auto#$i" FEDREC_X=1; done
run "ipc" FEDREC_SYMM=ipc
run "vmm-nonvls" FEDREC_SYMM=vmm FEDREC_NVLS=0
run "vmm-nvls" FEDREC_SYMM=vmm
EXTRA="--pipeline off" run "nopipe" FEDREC_X=1
