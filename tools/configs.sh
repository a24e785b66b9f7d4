# One line per BASELINE.json config (run through gpurun from the repo root); results in gpurun_out/<tag>/.
TAG=${1:-r01c}
O=gpurun_out/$TAG
mkdir -p $O
python bench.py --bindings 1000 --batch 100000 --steps 50 > $O/config1_topic_1k_100k.json 2> $O/config1.err
python bench.py --bindings 100000 --mode spec --steps 100 > $O/config2_topic_100k_spec.json 2> $O/config2s.err
python bench.py --bindings 100000 --mode compat --steps 100 > $O/config2_topic_100k_compat.json 2> $O/config2c.err
python bench_headers.py > $O/config3_headers_50k.json 2> $O/config3.err
python bench_decode.py > $O/config5_decode_1024seg.json 2> $O/config5a.err
python bench_decode.py --segments 1 > $O/config5_decode_1seg.json 2> $O/config5b.err
python bench_decode.py --route > $O/config5_decode_route.json 2> $O/config5c.err
python bench_fanout.py > $O/next_fanout.json 2> $O/fanout.err
python - $O <<'PY'
import json, sys, glob, os
for f in sorted(glob.glob(sys.argv[1] + "/*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(os.path.basename(f), "FAILED", e); continue
    cpu = d.get("cpu_baseline", {}).get("value")
    print("%-36s %10.1f M msgs/s  %7.3f ms/step  e2e %s  cpu(1 thread) %s" % (
        os.path.basename(f), d["value"] / 1e6, d["ms_per_step"],
        ("%.1f M" % (d["e2e"]["value"] / 1e6)) if "e2e" in d else "-", ("%.1f k" % (cpu / 1e3)) if cpu else "-"))
PY
