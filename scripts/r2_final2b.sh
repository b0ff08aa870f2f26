# Round 2 closing 2-GPU check after the macro sub-ops: sharded parity tests (four exchange variants) and the bench's
# sharded block under torchrun.
O=gpurun_out/${1:-r2w}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s > $O/dist_tests.log 2>&1; echo "dist tests exit $?"; tail -5 $O/dist_tests.log | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 2 --steps 1 --warmup 3 --skip-extras > $O/bench2.out 2> $O/bench2.err; echo "bench2 exit $?"
python - "$O" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1] + '/bench2.out').read().splitlines()[-1])
sh = d["sharded"]
print("value", d["value"], "e2e", d["e2e"]["value"], "parity", sh["parity"], "qv_strong", {k: sh["qv_strong"][k] for k in sh["qv_strong"] if k in ("gate_ms", "swaps", "fused_swaps", "nvlink_gbs_per_direction")})
PY
