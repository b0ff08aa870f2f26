# Round 2 closing check: the allocation-free entry point alone in a fresh process, then the driver's three
# steps (GPU tests with the new file collected last, smoke, bench).
O=gpurun_out/${1:-r2t}; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/no_alloc_alone.log 2>&1; echo "no_alloc alone exit $?" | tee -a $O/no_alloc_alone.log; tail -5 $O/no_alloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q --durations=8 > $O/tests.log 2>&1; echo "tests exit $?" | tee -a $O/tests.log; tail -14 $O/tests.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -2 $O/smoke.log
timeout 1200 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cut -c1-1500 $O/bench.json; tail -3 $O/bench.err
