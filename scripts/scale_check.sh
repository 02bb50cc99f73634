#!/bin/bash
# Multi-GPU health check for one world size: the driver's scaling bench command, several fresh process launches, chosen
# heap backends. Usage (under gpurun --gpus N):  bash scripts/scale_check.sh N [launches of the default config] [variants...]
# variants: ipc vmm-nonvls vmm-nvls nopipe collective. Writes gpurun_out/scale_check_N.txt (one summary line per launch;
# the stderr of a failed launch is kept in gpurun_out/scale_check_N_<label>.err).
N=${1:-4}; REP=${2:-3}; shift 2
OUT=gpurun_out/scale_check_${N}.txt
mkdir -p gpurun_out; : > $OUT
run() {   # label, extra bench args, env...
  label=$1; extra=$2; shift 2
  port=$((29600 + RANDOM % 300))
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $port bench.py --gpus $N --steps 20 --warmup 5 $extra > gpurun_out/sc.out 2> gpurun_out/sc.err
  rc=$?
  line=$(grep '"metric"' gpurun_out/sc.out | tail -1 | python -c 'import sys,json
for l in sys.stdin:
    d=json.loads(l); c=d.get("details") or d["config"]; print(json.dumps({"value":d["value"],"ms":d["ms_per_step"],"e2e":d["e2e"]["value"],"heap":c.get("peer_heap"),"transport":c.get("transport"),"opt":c.get("optimizer","")[:9],"degraded":c.get("degraded")}))')
  echo "N=$N $label rc=$rc $line" | tee -a $OUT
  if [ $rc -ne 0 ]; then
    cp gpurun_out/sc.err gpurun_out/scale_check_${N}_${label}.err
    grep -E "Error|mbarrier|fedrec|rank [0-9]" gpurun_out/sc.err | grep -v "^\s*$" | sort | uniq -c | sort -rn | head -12 | tee -a $OUT
  fi
}
for i in $(seq 1 $REP); do run "auto$i" "" FEDREC_X=1; done
for v in "$@"; do
  case $v in
    ipc) run ipc "" FEDREC_SYMM=ipc ;;
    vmm-nonvls) run vmm-nonvls "" FEDREC_SYMM=vmm FEDREC_NVLS=0 ;;
    vmm-nvls) run vmm-nvls "" FEDREC_SYMM=vmm ;;
    nopipe) run nopipe "--pipeline off" FEDREC_X=1 ;;
    collective) run collective "--transport collective" FEDREC_X=1 ;;
  esac
done
