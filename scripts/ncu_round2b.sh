# Round-2 closing profile (after the macro sub-ops): (1) plain bench, (2) ncu launch list of the SAME command,
# (3) --set full over one SU2 evaluation's fused passes.  Numbers printed under ncu are never bench values.
O=gpurun_out/${1:-ncu_r2b}; mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --skip-sharded"
timeout 900 $CMD > $O/bench_plain.json 2> $O/bench_plain.err; echo "plain bench exit $?"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench.csv \
  $CMD > $O/bench_under_ncu.log 2>&1; echo "ncu launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused_pass -s 12 -c 12 \
  -o $O/su2_fused -f python scripts/profile_run.py su2 > $O/ncu_su2.log 2>&1; echo "ncu fused exit $?"
for r in su2_fused; do
  if [ -f $O/$r.ncu-rep ]; then
    python scripts/ncu_summary.py $O/$r.ncu-rep > $O/$r.summary.txt 2>&1
    ncu -i $O/$r.ncu-rep --page raw --csv > $O/$r.raw.csv 2>/dev/null
    rm -f $O/$r.ncu-rep
  fi
done
ls -la $O
