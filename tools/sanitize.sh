# compute-sanitizer over the small-shape GPU tests (run through gpurun from the repo root): memcheck and racecheck
# logs land in gpurun_out/<tag>/ and are copied to profiles/ by hand.  racecheck matters here: k_offsets is a decoupled
# look-back over volatile state, k_write's warps share per-warp buffers across paths, the fan-out sorts use shared memory.
TAG=${1:-san}
O=gpurun_out/$TAG
mkdir -p $O
SEL='test_triematcher_main_kat or test_compat_edge_vectors or test_random_small_tables or test_direct_and_fanout or test_runs_of_dots or test_transpose_edge_cases or test_static_scratch_entry_point_refuses'
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 --log-file $O/sanitizer_$tool.log \
    python -m pytest tests/test_gpu_route.py tests/test_fanout.py tests/test_zz_no_alloc.py -m gpu -x -q -k "$SEL" > $O/sanitizer_${tool}_pytest.log 2>&1
  echo "$tool exit $?" >> $O/sanitizer_$tool.log
  tail -3 $O/sanitizer_$tool.log
done
