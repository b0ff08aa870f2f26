# Round check: GPU parity tests, smoke, both bench arms, short profile runs.
O=gpurun_out/${1:-chk}; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; echo "tests exit $?" | tee -a $O/tests.log; tail -3 $O/tests.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -2 $O/smoke.log
timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-200 | tee $O/su2.txt
timeout 300 python scripts/profile_run.py sampler 2>&1 | cut -c1-200 | tee $O/sampler.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cat $O/bench.json
if [ "${2:-ref}" = "ref" ]; then timeout 900 python bench.py --impl reference > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"; cut -c1-300 $O/bench_ref.json; fi
