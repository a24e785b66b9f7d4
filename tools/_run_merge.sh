mkdir -p gpurun_out/r02u
python -m pytest tests/test_sharded.py tests/test_gpu_route.py -x -q -m gpu -k "slices_on_one_gpu or several_long_lists or random_small" 2>&1 | tail -4
for g in 8 4 2; do python tools/bench_merge.py $g 20 2>&1 | tail -2 | head -1; done
