# Round 2, 2-GPU call: exchange variants.  default = swaps ride inside fused passes (scatter, else gather), standalone
# exchanges are gate-less scatter passes; PUSH=0: standalone exchanges through k_swap_p2p; FUSE=0: round-1 behaviour.
O=gpurun_out/${1:-r2f}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s > $O/dist_tests.log 2>&1; echo "dist tests exit $?"; tail -3 $O/dist_tests.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bilinear" > $O/bilinear_test.log 2>&1; echo "bilinear test exit $?"; tail -2 $O/bilinear_test.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 scripts/config_runs.py $2 $3 $4 2>&1 | grep -v "^\[fused\]\|^ *->\|^pass\|^\*\*\*\|OMP_NUM\|^$\|tabs\|^ *(fused" | tail -8; }
for cfg in "qv 33 16" "tfim 33 4" "qv 31 16"; do
  set -- $cfg
  echo "== $cfg default"; run 29655 $1 $2 $3 | tee $O/$1$2_default.log
  echo "== $cfg PUSH=0"; QSV_DIST_PUSH_SWAP=0 run 29656 $1 $2 $3 | tee $O/$1$2_nopush.log
  echo "== $cfg FUSE=0"; QSV_DIST_FUSE_SWAP=0 run 29657 $1 $2 $3 | tee $O/$1$2_unfused.log
done
echo "== tfim 33 4 default, traced"; QSV_TRACE=1 run 29658 tfim 33 4 | grep "swap" | tee $O/tfim33_trace.log
