import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1:], "value %.3e ms/step %.3f"%(d["value"], d["ms_per_step"]), {k: round(v,3) for k,v in d["kernel_ms"].items()}, "lm_it/s %.1f pcg/it %.0f e2e %.3e"%(d["lm_iters_per_sec"], d["pcg_iters_per_lm_iter"], d["e2e"]["value"]), {k: round(v, 3) for k, v in d.get("lm_phase_ms", {}).items()})
