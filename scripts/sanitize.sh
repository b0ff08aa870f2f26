# compute-sanitizer over the small-n kernel matrix: memcheck + racecheck logs under $O
O=gpurun_out/${1:-sanitize}; mkdir -p $O
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 python scripts/sanitize_matrix.py > $O/sanitizer_$tool.log 2>&1
  echo "$tool exit $?" | tee -a $O/sanitizer_$tool.log; tail -5 $O/sanitizer_$tool.log
done
