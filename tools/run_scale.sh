#!/bin/bash
# Dev: every bench workload at N GPUs of one box.  usage: tools/run_scale.sh N [epochs]   (logs -> gpurun_out/scale_N/)
N=${1:-1}; EP=${2:-200}
OUT=gpurun_out/scale_$N; mkdir -p $OUT
ONLY=${ONLY:-"infer train train64 cv preprocess stress"}
run() { # name, args...
  local name=$1; shift
  case " $ONLY " in *" $name "*) ;; *) return;; esac
  if [ "$N" = "1" ]; then timeout 900 python bench.py --gpus 1 "$@" > $OUT/$name.log 2>&1
  else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N "$@" > $OUT/$name.log 2>&1; fi
  tail -1 $OUT/$name.log | cut -c1-220
}
run infer --no-extras
run train --workload train --no-extras
run train64 --workload train --no-extras --batch 64
run cv --workload cv --epochs $EP
run preprocess --workload preprocess
run stress --workload stress --steps 10 --warmup 3
