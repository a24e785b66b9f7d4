# usage (through gpurun): bash tools/try_variants.sh variants/*.so   — bench every prebuilt library variant
cp chanamq_b200/libchanagpu.so /tmp/lib_keep.so
for v in "$@"; do
  cp $v chanamq_b200/libchanagpu.so
  for md in ${MODES:-compat}; do
    python bench.py --no-cpu-baseline --steps ${STEPS:-50} --mode $md > /tmp/var.json 2> /tmp/var.err || { echo "$v $md FAILED"; tail -3 /tmp/var.err; continue; }
    python - "$v" $md <<'PY'
import json, sys
d = json.loads(open("/tmp/var.json").read().strip().splitlines()[-1])
k = d["roofline"]["per_kernel_ms"]
print("%-34s %-6s %.3f G msgs/s  step %.3f ms  tok %.3f probe %.3f off %.3f write %.3f rest %.3f  e2e %.1f M" % (
    sys.argv[1], sys.argv[2], d["value"] / 1e9, d["ms_per_step"], k["k_tokenize"], k["k_probe"], k["k_offsets"], k["k_write"], k.get("k_write_rest", 0), d["e2e"]["value"] / 1e6))
PY
  done
done
cp /tmp/lib_keep.so chanamq_b200/libchanagpu.so
