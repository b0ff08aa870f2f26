# ncu --set full over the 13 launches of the second 388-gate SU2 iteration (per-pass view of k_fused_pass)
O=gpurun_out/${1:-ncu_su2}; mkdir -p $O
V=${2:-0}
QSV_FUSED3=$V timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused_pass -s 12 -c 12 \
  -o $O/su2_f$V -f python scripts/profile_run.py su2 > $O/ncu_su2_f$V.log 2>&1
tail -5 $O/ncu_su2_f$V.log
ls -la $O
