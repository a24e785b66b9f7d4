# usage: bash tools/variants.sh "<defs A>" "<defs B>" ...   — rebuild with each set of -D switches and bench both modes
for defs in "$@"; do
  CHQ_NVCC_DEFS="$defs" python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1 || { echo "build failed: $defs"; continue; }
  for md in compat spec; do
    python bench.py --no-cpu-baseline --steps 50 --mode $md > gpurun_out/var.json 2> gpurun_out/var.err
    python - "$defs" $md <<'PY'
import json, sys
d = json.loads(open("gpurun_out/var.json").read().strip().splitlines()[-1])
k = d["roofline"]["per_kernel_ms"]
print("%-40s %-6s %.3f G msgs/s  step %.3f ms  tok %.3f probe %.3f off %.3f write %.3f  e2e %.1f M" % (
    sys.argv[1] or "(default)", sys.argv[2], d["value"] / 1e9, d["ms_per_step"], k["k_tokenize"], k["k_probe"], k["k_offsets"], k["k_write"], d["e2e"]["value"] / 1e6))
PY
  done
done
