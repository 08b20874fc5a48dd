python -m pytest tests -m gpu -x -q -k "filter or resample or sosfilt" 2>&1 | tail -3
python bench.py --workload preprocess --no-extras 2>&1 | tail -1 | python -c "import json,sys;d=json.loads(sys.stdin.read());print(d['value'],d['roofline']['ms_per_launch'],d['roofline']['two_kernel_chain_ms'])"
