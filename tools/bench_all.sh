#!/bin/bash
# every BASELINE config at N GPUs (N = $1, default 1): JSON lines into gpurun_out/r02_bench_<config>_n<N>.json
N=${1:-1}; shift
CFGS=${*:-c2 nst c3_4k c4_reconet300 c5_region}
mkdir -p gpurun_out
for c in $CFGS; do
  out=gpurun_out/r02_bench_${c}_n${N}.json
  if [ "$N" = "1" ]; then
    timeout 900 python bench.py --config $c --steps ${STEPS:-3} --warmup 3 > $out 2> gpurun_out/bench_${c}_n${N}.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --config $c --gpus $N --steps ${STEPS:-3} --warmup 3 > $out 2> gpurun_out/bench_${c}_n${N}.err
  fi
  echo "$c n=$N rc=$?"; python - "$out" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    r = d["roofline"]
    print("  value %.1f %s | e2e %.1f | ms/step %.2f | roofline %s %.1f/%.1f %s = %.3f | cpu %s | clocks %s" % (d["value"], d["unit"], d["e2e"]["value"], d["ms_per_step"], r["bound"], r["achieved"], r["peak"], r["unit"], r["frac"], (d.get("cpu_baseline") or {}).get("value"), d["clocks"]))
except Exception as e:
    print("  (no JSON line)", e); print(open(sys.argv[1].replace("r02_bench_", "bench_").replace(".json", ".err")).read()[-1500:])
PY
done
