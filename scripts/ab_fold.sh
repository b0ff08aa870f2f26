# A/B: product-state prologue folded into the first fused pass (QSV_PRODUCT_FOLD=0 disables it)
O=gpurun_out/${1:-abfold}; mkdir -p $O
if [ "${2:-tests}" = "tests" ]; then timeout 1400 python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; tail -3 $O/tests.log; fi
echo "== fold on"; timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-160 | tee $O/on.txt
echo "== fold off"; QSV_PRODUCT_FOLD=0 timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-160 | tee $O/off.txt
QSV_TRACE=1 timeout 300 python scripts/trace_run.py su2 28 2>&1 | grep -A1 "no read" | head -6
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cut -c1-1200 $O/bench.json
