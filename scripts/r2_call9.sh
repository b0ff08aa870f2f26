# Round 2: shared-prefix caching + SparseObservable inputs on the device, then the bench with its shared_prefix block.
O=gpurun_out/${1:-r2u}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_prefix.py tests/test_gpu_observables.py tests/test_zz_no_alloc.py -m gpu -x -q > $O/tests_new.log 2>&1; echo "new tests exit $?" | tee -a $O/tests_new.log; tail -15 $O/tests_new.log
timeout 1200 python bench.py --skip-sharded > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -3 $O/bench.err
python - <<'PY'
import json,sys
d=json.loads(open(sys.argv[1] if len(sys.argv)>1 else 'gpurun_out/r2u/bench.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value']); print(json.dumps(d.get('shared_prefix'), indent=1))
PY
