# Round 2, 2-GPU call: fused swap pass with staged prefetch (A/B against no staging and against the standalone swap).
O=gpurun_out/${1:-r2c}; mkdir -p $O
timeout 600 python -m pytest tests/test_adapter.py tests/test_gpu_large.py -m gpu -x -q -s > $O/tests_new.log 2>&1; echo "new tests exit $?"; tail -12 $O/tests_new.log
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s > $O/dist_tests.log 2>&1; echo "dist tests exit $?"; tail -4 $O/dist_tests.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 scripts/config_runs.py qv $2 $3 2>&1 | grep -v "^\[fused\]\|^ *->\|^pass\|^\*\*\*\|OMP_NUM\|^$" | tail -12; }
echo "== qv31 d8, staged (default)"; QSV_TRACE=1 run 29655 31 8 | tee $O/qv31_staged.log
echo "== qv31 d8, no staging"; QSV_FUSED_SWAP_STAGE=0 QSV_TRACE=1 run 29656 31 8 | tee $O/qv31_nostage.log
echo "== qv31 d8, standalone swap"; QSV_DIST_FUSE_SWAP=0 QSV_TRACE=1 run 29657 31 8 | tee $O/qv31_unfused.log
echo "== qv33 d16 staged"; run 29658 33 16 | tee $O/qv33_staged.log
echo "== qv33 d16 standalone"; QSV_DIST_FUSE_SWAP=0 run 29659 33 16 | tee $O/qv33_unfused.log
