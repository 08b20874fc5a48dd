ncu --set full --clock-control none --import-source on -k regex:'transformer_(fwd|bwd)' -s 8 -c 2 -o gpurun_out/r01_tr_b64 --force-overwrite python bench.py --workload train --batch 64 --no-extras --steps 2 --warmup 3 > gpurun_out/ncu_tr.log 2>&1
ls -la gpurun_out/r01_tr_b64.ncu-rep
