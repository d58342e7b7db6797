#!/bin/bash
# micro-batch sweep of the bench (L2 residency of the trunk activations vs launch count): tools/mb_sweep.sh "2 3 4 8"
mkdir -p gpurun_out
for mb in ${1:-2 3 4 6 8}; do
  NST_MICROBATCH=$mb timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/mb_$mb.json 2> gpurun_out/mb_$mb.err
  python - $mb <<'PY'
import json, sys
mb = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/mb_{mb}.json") if l.startswith("{")][-1])
    print("mb", mb, "value %.1f e2e %.1f" % (d["value"], d["e2e"]["value"]), d["clocks"], "trunk frac", d["roofline"]["frac"])
    print("   ", " ".join("%s %.4f" % (l["name"], l["ms_per_frame"]) for l in d["layers"]))
except Exception as e:
    print("mb", mb, "failed", e); print(open(f"gpurun_out/mb_{mb}.err").read()[-800:])
PY
done
