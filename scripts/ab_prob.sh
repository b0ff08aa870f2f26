# A/B: probability-only store of the last basis-change pass (QSV_NO_PROB_STORE=1 disables it)
O=gpurun_out/${1:-abprob}; mkdir -p $O
timeout 1400 python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; tail -3 $O/tests.log
echo "== prob store on"; timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-160 | tee $O/on.txt
echo "== prob store off"; QSV_NO_PROB_STORE=1 timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-160 | tee $O/off.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cut -c1-900 $O/bench.json
