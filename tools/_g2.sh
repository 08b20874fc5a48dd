python -m pytest tests -m gpu -x -q 2>&1 | tail -2
bash tools/run_scale.sh 1 200
python bench.py > gpurun_out/bench_default.log 2>&1; tail -1 gpurun_out/bench_default.log | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log | cut -c1-300
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_train_b64.csv python bench.py --workload train --batch 64 --steps 2 --warmup 3 --no-extras > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_preprocess.csv python bench.py --workload preprocess --steps 2 --warmup 3 --no-extras > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'sosfiltfilt_resample' -s 4 -c 1 -o gpurun_out/r01_fused --force-overwrite python bench.py --workload preprocess --no-extras --steps 2 --warmup 3 > /dev/null 2>&1
ls gpurun_out
