#!/bin/bash
# tile-size sweep of the fused linearisation kernel (run under gpurun)
for wl in C D; do
  for t in 320 640 960 1312; do
    B200BA_TILE_OBS=$t python bench.py --workload $wl --steps 10 --no-cpu-baseline --e2e-steps 1 > gpurun_out/sweep_${wl}_${t}.log 2>&1
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/sweep_${wl}_${t}.log").read().strip().splitlines()[-1])
    print("${wl}", ${t}, "obs/s %.3e"%d["value"], "ms/step %.3f"%d["ms_per_step"], d["kernel_ms"], d["tiles"], "fp64 frac %.3f peak %.1f"%(d["roofline_fp64"]["frac"], d["roofline_fp64"]["peak"]), "lm it/s %.1f"%d["lm_iters_per_sec"])
except Exception as e:
    print("${wl}", ${t}, "FAILED", e); print(open("gpurun_out/sweep_${wl}_${t}.log").read()[-800:])
PY
  done
done
