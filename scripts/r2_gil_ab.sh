# A/B of the GIL hand-off interval for the two host lanes: the e2e arm and the shared_prefix block of bench.py
O=gpurun_out/${1:-r2x}; mkdir -p $O
for v in 0 0.0002; do
  QSV_SWITCH_INTERVAL=$v timeout 600 python bench.py --skip-sharded > $O/bench_si_$v.json 2> $O/bench_si_$v.err; echo "interval $v exit $?"
  python - "$O/bench_si_$v.json" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().splitlines()[-1])
sp = d["shared_prefix"]
print("value %.2f e2e %.2f | shared_prefix off %.2f on %.2f ms/eval | config0 %s" % (d["value"], d["e2e"]["value"], sp["cache_off_ms_per_eval"], sp["cache_on_ms_per_eval"], {k: v for k, v in d["config0"].items() if "ms" in k and isinstance(v, float)}))
PY
done
