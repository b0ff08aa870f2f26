# A/B of the second-wave start delay of k_fused_pass (QSV_STAGGER_NS = ns per phase)
O=gpurun_out/${1:-abstag}; mkdir -p $O
for v in 0 600 1200 2000 3000; do
  echo "== QSV_STAGGER_NS=$v"
  QSV_STAGGER_NS=$v timeout 300 python scripts/profile_run.py su2 2>&1 | grep "^\[1\]" | cut -c1-40 | tee $O/su2_$v.txt
  QSV_STAGGER_NS=$v timeout 300 python scripts/profile_run.py stream 2>&1 | grep "^\[1\] QV" | cut -c1-60 | tee $O/stream_$v.txt
  QSV_STAGGER_NS=$v QSV_TRACE=1 timeout 300 python scripts/trace_run.py qv 30 4 > $O/trace_qv_$v.txt 2>&1
  QSV_STAGGER_NS=$v QSV_TRACE=1 timeout 300 python scripts/trace_run.py su2 28 > $O/trace_su2_$v.txt 2>&1
done
