# A/B of the lifted plane rotations (QSV_NO_LIFT=1 disables them)
O=gpurun_out/${1:-ablift}; mkdir -p $O
for v in off on; do
  if [ $v = off ]; then export QSV_NO_LIFT=1; else unset QSV_NO_LIFT; fi
  for w in "tfim 30 10" "u3 30 4" "su2 28"; do
    set -- $w
    timeout 300 python scripts/trace_run.py $w > $O/plain_${1}_$v.txt 2>&1; echo "$v $w: $(grep total $O/plain_${1}_$v.txt)"
  done
  timeout 300 python scripts/profile_run.py su2 2>&1 | grep "^\[1\]" | cut -c1-40 | tee $O/su2_$v.txt
done
unset QSV_NO_LIFT
QSV_TRACE=1 timeout 300 python scripts/trace_run.py su2 28 > $O/trace_su2_on.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_density.py tests/test_basic_simulator.py -m gpu -x -q 2>&1 | tail -3
