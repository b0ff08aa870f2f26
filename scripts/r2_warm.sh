# descriptor warm-up for small states: A/B at 20 q (per-pass kernel time under QSV_TRACE, untraced stream time)
O=gpurun_out/${1:-r2i}; mkdir -p $O
for v in 0 1; do
  echo "== QSV_NO_WARM=$v traced"; QSV_NO_WARM=$v QSV_TRACE=1 timeout 300 python scripts/profile_run.py qv 20 2>&1 | grep "\->" | tail -6
  echo "== QSV_NO_WARM=$v"; QSV_NO_WARM=$v timeout 300 python scripts/profile_run.py qv 20 2>&1 | grep "^\[" | cut -c1-60
  QSV_NO_WARM=$v timeout 300 python scripts/profile_run.py qv 22 2>&1 | grep "^\[1" | cut -c1-60
done | tee $O/warm_ab.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_density.py -m gpu -x -q > $O/tests.log 2>&1; echo "tests exit $?"; tail -2 $O/tests.log
python - <<'PY' | tee -a gpurun_out/r2i/warm_ab.txt
import sys, json, time
sys.path.insert(0, '/root/repo')
import bench
for it in range(2):
    print(json.dumps({k: v for k, v in bench.config0_block(0, with_cpu=False).items() if k not in ("workload", "note")}))
PY
