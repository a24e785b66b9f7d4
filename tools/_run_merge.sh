python -m pytest tests/test_sharded.py tests/test_gpu_route.py -x -q -m gpu -k "slices_on_one_gpu or several_long_lists" 2>&1 | tail -3
for g in 8 4 2; do python tools/bench_merge.py $g 20 2>&1 | tail -2 | head -1; done
