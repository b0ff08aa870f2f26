# Round 2, 1-GPU call: bilinear phase tables (A/B), full GPU tests, sanitizer matrix.
O=gpurun_out/${1:-r2e}; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -x -q > $O/tests.log 2>&1; echo "tests exit $?" | tee -a $O/tests.log; tail -4 $O/tests.log
for m in qft tfim; do
  echo "== $m bilinear"; timeout 300 python scripts/profile_run.py $m 2>&1 | cut -c1-220 | tee $O/${m}_bilinear.txt
  echo "== $m generic tables"; QSV_NO_BILINEAR=1 timeout 300 python scripts/profile_run.py $m 2>&1 | cut -c1-220 | tee $O/${m}_generic.txt
done
echo "== su2"; timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-220 | tee $O/su2.txt
echo "== qv"; timeout 300 python scripts/profile_run.py qv 2>&1 | cut -c1-220 | tee $O/qv.txt
bash scripts/sanitize.sh $(basename $O)
