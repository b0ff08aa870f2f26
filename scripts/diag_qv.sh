# QV fused-pass diagnosis: L2-resident (22 q) vs HBM (30 q) per-pass times, then ncu --set full on three 30 q passes
O=gpurun_out/${1:-diagqv}; mkdir -p $O
for n in 22 24 30; do
  QSV_TRACE=1 timeout 300 python scripts/trace_run.py qv $n 4 > $O/trace_qv_$n.txt 2>&1
  QSV_TRACE=1 timeout 300 python scripts/trace_run.py su2 $n > $O/trace_su2_$n.txt 2>&1
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused_pass -s 10 -c 3 \
  -o $O/qv30 -f python scripts/trace_run.py qv 30 4 > $O/ncu_qv30.log 2>&1
tail -3 $O/ncu_qv30.log; ls -la $O
