# Round 2, final 1-GPU check: the driver's three steps (GPU tests, smoke, bench) + the 20 q per-pass trace.
O=gpurun_out/${1:-r2h}; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -x -q --durations=8 > $O/tests.log 2>&1; echo "tests exit $?" | tee -a $O/tests.log; tail -14 $O/tests.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -2 $O/smoke.log
timeout 1200 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cut -c1-1500 $O/bench.json; tail -3 $O/bench.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"; cut -c1-700 $O/bench_ref.json
QSV_TRACE=1 timeout 300 python scripts/profile_run.py qv 20 > $O/qv20_trace.txt 2>&1; grep -c "\->" $O/qv20_trace.txt; grep "\->" $O/qv20_trace.txt | tail -12; grep "^\[1\]" $O/qv20_trace.txt | cut -c1-160
timeout 300 python scripts/sanitize_matrix.py > $O/matrix.log 2>&1; tail -1 $O/matrix.log
