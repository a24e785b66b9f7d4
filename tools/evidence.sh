# Collects the round's evidence on one B200 (run through gpurun from the repo root): GPU tests, the bench lines,
# the ncu launch list and one full capture of each route kernel.  Every ncu pass runs only after the same
# command has exited 0 without ncu.  Results land in gpurun_out/<tag>/; tools/refresh_profiles.py copies them to profiles/.
set -x
TAG=${1:-r02}
O=gpurun_out/$TAG
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python bench.py > $O/bench.json 2> $O/bench.err || exit 1
python bench.py --impl reference > $O/bench_ref.json 2> $O/bench_ref.err
python bench.py --mode spec --no-cpu-baseline --no-configs > $O/bench_spec.json 2> $O/bench_spec.err
python bench_fanout.py > $O/bench_fanout.json 2> $O/bench_fanout.err
python bench_encode.py > $O/bench_encode.json 2> $O/bench_encode.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-configs > $O/plain.log 2>&1 || exit 2
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-configs > $O/ncu_l.log 2>&1
for k in write probe tokenize; do
  ncu --set full --clock-control none --import-source on -k regex:k_$k\$ --launch-skip 2 -c 1 -o $O/prof_$k python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-configs > $O/ncu_$k.log 2>&1
done
tail -3 $O/pytest_gpu.log; cat $O/bench.json
