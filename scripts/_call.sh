set -x
O=gpurun_out/r1b; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/gpu_tests.log 2>&1; tail -15 $O/gpu_tests.log
for tb in 12 11; do QSV_TILE_BITS=$tb timeout 300 python scripts/profile_run.py su2 > $O/su2_T$tb.log 2>&1; cat $O/su2_T$tb.log; done
QSV_USE_CLUSTER=0 timeout 300 python scripts/profile_run.py su2 > $O/su2_legacy.log 2>&1; cat $O/su2_legacy.log
for tb in 12 11; do QSV_TILE_BITS=$tb timeout 300 python scripts/profile_run.py stream > $O/stream_T$tb.log 2>&1; cat $O/stream_T$tb.log; done
