# A/B of the complex-1q -> diag . real . diag split (QSV_NO_ZYZ=1 disables it)
O=gpurun_out/${1:-abzyz}; mkdir -p $O
for v in off on; do
  if [ $v = off ]; then export QSV_NO_ZYZ=1; else unset QSV_NO_ZYZ; fi
  for w in "tfim 30 10" "u3 30 4" "su2 28" "qft 30"; do
    set -- $w
    timeout 300 python scripts/trace_run.py $w > $O/plain_${1}_$v.txt 2>&1; echo "$v $w: $(grep total $O/plain_${1}_$v.txt)"
    QSV_TRACE=1 timeout 300 python scripts/trace_run.py $w > $O/trace_${1}_$v.txt 2>&1
  done
done
unset QSV_NO_ZYZ
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_density.py -m gpu -x -q 2>&1 | tail -3
