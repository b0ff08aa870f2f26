O=gpurun_out/${1:-abfold}; mkdir -p $O
echo "== fold on"; timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-160 | tee $O/on.txt
echo "== fold off"; QSV_PRODUCT_FOLD=0 timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-160 | tee $O/off.txt
QSV_TRACE=1 timeout 300 python scripts/trace_run.py su2 28 2>&1 | grep -A1 generated | head -6
