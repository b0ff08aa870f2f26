# kernel experiment: timings of the four benchmark streams + parity subset
O=gpurun_out/${1:-r2m}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/tests.log 2>&1; echo "tests exit $?"; tail -2 $O/tests.log
for m in su2 qv qft tfim; do timeout 300 python scripts/profile_run.py $m 2>&1 | grep "^\[1\]" | cut -c1-70; done | tee $O/timings.txt
