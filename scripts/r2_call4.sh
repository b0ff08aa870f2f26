# Round 2, 2-GPU call: fused swap pass, scatter (push) vs gather (pull) vs standalone exchange.
O=gpurun_out/${1:-r2d}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s > $O/dist_tests.log 2>&1; echo "dist tests exit $?"; grep -c "scatter\] amp" $O/dist_tests.log; grep "scatter\]" $O/dist_tests.log | head -12; tail -4 $O/dist_tests.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 scripts/config_runs.py qv $2 $3 2>&1 | grep -v "^\[fused\]\|^ *->\|^pass\|^\*\*\*\|OMP_NUM\|^$\|tabs" | tail -12; }
echo "== qv31 d8, scatter+gather (default)"; QSV_TRACE=1 run 29655 31 8 | tee $O/qv31_scatter.log
echo "== qv31 d8, gather only"; QSV_DIST_FUSE_SWAP=1 QSV_TRACE=1 run 29656 31 8 | tee $O/qv31_gather.log
echo "== qv31 d8, standalone swap"; QSV_DIST_FUSE_SWAP=0 QSV_TRACE=1 run 29657 31 8 | tee $O/qv31_unfused.log
echo "== qv33 d16 scatter+gather"; run 29658 33 16 | tee $O/qv33_scatter.log
echo "== qv33 d16 standalone"; QSV_DIST_FUSE_SWAP=0 run 29659 33 16 | tee $O/qv33_unfused.log
