# kernel experiment: timings of the four benchmark streams + parity subset (QSV_NO_D1L4=1 = without the macro sub-ops)
O=gpurun_out/${1:-r2m}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_density.py tests/test_backend_primitives.py -m gpu -x -q > $O/tests.log 2>&1; echo "tests exit $?"; tail -2 $O/tests.log
for v in 0 1; do echo "== QSV_NO_D1L4=$v"; for m in su2 qv qft tfim; do QSV_NO_D1L4=$v timeout 300 python scripts/profile_run.py $m 2>&1 | grep "^\[1\]" | cut -c1-70; done; done | tee $O/timings.txt
