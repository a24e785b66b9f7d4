# compute-sanitizer (memcheck) over the quick GPU parity tests; run through gpurun from the repo root, in a call
# of its own (never together with ncu):   gpurun --timeout 1500 -- 'bash tools/sanitize.sh'
set -x
O=gpurun_out/sanitize
mkdir -p $O
SAN="compute-sanitizer --tool memcheck --error-exitcode 77 --print-limit 20"
run() {  # name, pytest args...
  local name=$1; shift
  timeout 900 $SAN python -m pytest "$@" -m gpu -x -q > $O/$name.log 2>&1
  echo "$name rc=$?" >> $O/summary.txt
  grep -E "ERROR SUMMARY|passed|failed" $O/$name.log | tail -3 >> $O/summary.txt
}
: > $O/summary.txt
run route tests/test_gpu_route.py -k "kat or edge_vectors or longer_than or unbind or empty_table or word_lengths or several_long or direct_and or tile_edges"
run fanout tests/test_fanout.py -k "against_the_oracle or edge_cases"
run headers tests/test_headers.py -k "rule_examples or compat_mode"
run decode tests/test_gpu_decode.py -k "reference_vectors or empty_inputs or chunked_scan"
run noalloc tests/test_zz_no_alloc.py
cat $O/summary.txt
