set -x
mkdir -p gpurun_out/r01s
O=gpurun_out/r01s
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python bench.py > $O/bench.json 2> $O/bench.err || exit 1
python bench.py --impl reference > $O/bench_ref.json 2> $O/bench_ref.err
python bench.py --mode spec --no-cpu-baseline > $O/bench_spec.json 2> $O/bench_spec.err
python bench_decode.py > $O/bench_decode.json 2> $O/bench_decode.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/plain.log 2>&1 || exit 2
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_write --launch-skip 2 -c 1 -o $O/prof_write python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_w.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_probe --launch-skip 2 -c 1 -o $O/prof_probe python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_p.log 2>&1
tail -3 $O/pytest_gpu.log; cat $O/bench.json
