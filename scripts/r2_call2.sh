# Round 2, 2-GPU call: sharded parity (fused swap pass / standalone p2p / nccl) and the bench's sharded block at N = 2.
O=gpurun_out/${1:-r2b}; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s > $O/dist_tests.log 2>&1; echo "dist tests exit $?"; tail -40 $O/dist_tests.log
QSV_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 scripts/config_runs.py qv 31 8 > $O/qv31_trace.log 2>&1; grep -v "^\[fused\]\|^ *->\|^pass" $O/qv31_trace.log | tail -30
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29656 bench.py --gpus 2 --steps 1 --warmup 1 --skip-extras > $O/bench2.json 2> $O/bench2.err; echo "bench exit $?"; python - <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1] if len(sys.argv)>1 else '/root/repo/gpurun_out/'+'r2b/bench2.json') if l.startswith('{')][-1])
    print(json.dumps(d.get('sharded'), indent=1)[:5000])
except Exception as e:
    print('no json', e)
PY
grep -i "nranks\|error\|Traceback" $O/bench2.err | head -20
