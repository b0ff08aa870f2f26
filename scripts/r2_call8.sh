# Round 2, 8-GPU call: world-8 sharded parity, the driver's N = 8 bench command, exchange A/B at 33 q / 36 q.
O=gpurun_out/${1:-r2g8}; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s -k "8" > $O/dist8.log 2>&1; echo "dist world-8 exit $?"; grep -c "amp err" $O/dist8.log; tail -3 $O/dist8.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 8 --steps 2 --warmup 3 > $O/bench8.json 2> $O/bench8.err; echo "bench8 exit $?"
python - <<'PY'
import json
try:
    d = json.loads([l for l in open('/root/repo/gpurun_out/r2g8/bench8.json') if l.startswith('{')][-1])
    print("value", d["value"], "e2e", d["e2e"]["value"])
    print(json.dumps(d.get("sharded"), indent=1)[:7000])
except Exception as e:
    print("no json", e)
PY
grep -i "nranks" $O/bench8.err | head -3; grep -i "error\|Traceback" $O/bench8.err | head
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 scripts/config_runs.py $2 $3 $4 2>&1 | grep -v "^\[fused\]\|^ *->\|^pass\|^\*\*\*\|OMP_NUM\|^$\|tabs\|^ *(fused\|NCCL version" | tail -14; }
echo "== qv 33 16 default (traced swaps)"; QSV_TRACE=1 run 29711 qv 33 16 | tee $O/qv33_default_traced.log
echo "== qv 33 16 default"; run 29712 qv 33 16 | tee $O/qv33_default.log
echo "== qv 33 16 PUSH=0"; QSV_DIST_PUSH_SWAP=0 run 29713 qv 33 16 | tee $O/qv33_nopush.log
echo "== qv 33 16 FUSE=0"; QSV_DIST_FUSE_SWAP=0 run 29714 qv 33 16 | tee $O/qv33_unfused.log
echo "== qv 36 12 FUSE=0"; QSV_DIST_FUSE_SWAP=0 run 29715 qv 36 12 | tee $O/qv36_unfused.log
