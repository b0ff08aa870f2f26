# Round 2: after the host-side lowering speed-ups (constant gate matrices shared, plain attribute copy of operations)
O=gpurun_out/${1:-r2z}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; echo "tests exit $?"; tail -3 $O/tests.log
timeout 600 python bench.py --skip-sharded > $O/bench.json 2> $O/bench.err; echo "bench exit $?"
python - "$O/bench.json" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().splitlines()[-1])
sp = d["shared_prefix"]
print("value %.2f e2e %.2f | shared_prefix off %.2f on %.2f ms/eval | config0 %s" % (d["value"], d["e2e"]["value"], sp["cache_off_ms_per_eval"], sp["cache_on_ms_per_eval"], {k: v for k, v in d["config0"].items() if isinstance(v, float)}))
PY
