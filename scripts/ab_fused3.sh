# A/B of k_fused_pass (2 CTAs/SM, one buffer each) against k_fused_pass3 (1 CTA/SM, 2 groups, 3 buffers)
O=gpurun_out/${1:-ab3}; mkdir -p $O
QSV_FUSED3=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/tests_fused3.log 2>&1; echo "tests(fused3) exit $?"; tail -3 $O/tests_fused3.log
for v in 0 1; do
  echo "== QSV_FUSED3=$v"
  QSV_FUSED3=$v timeout 300 python scripts/profile_run.py su2 2>&1 | tee $O/su2_f$v.txt
  QSV_FUSED3=$v timeout 300 python scripts/profile_run.py stream 2>&1 | tee $O/stream_f$v.txt
  QSV_FUSED3=$v QSV_TRACE=1 timeout 300 python scripts/trace_run.py su2 28 > $O/trace_su2_f$v.txt 2>&1
  QSV_FUSED3=$v QSV_TRACE=1 timeout 300 python scripts/trace_run.py qv 30 4 > $O/trace_qv_f$v.txt 2>&1
done
