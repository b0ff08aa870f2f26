# Round 2, first GPU call: full GPU test suite (durations), smoke, the new bench blocks at N = 1.
O=gpurun_out/${1:-r2a}; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,memory.free --format=csv > $O/gpu.txt 2>&1
free -g > $O/host.txt; nproc >> $O/host.txt
timeout 1800 python -m pytest tests -m gpu -x -q --durations=15 > $O/tests.log 2>&1; echo "tests exit $?" | tee -a $O/tests.log; tail -25 $O/tests.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -2 $O/smoke.log
timeout 1200 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cut -c1-6000 $O/bench.json; tail -5 $O/bench.err
