# 8-GPU round check: cfg 5 (36 q TFIM), 36 q QV depth 12, the 33 q QV strong-scaling point, bench at N = 8
O=gpurun_out/${1:-r1n8}; mkdir -p $O
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 "$@"; }
run scripts/config_runs.py tfim 36 10 2>$O/tfim36.err | tee $O/tfim36.json
run scripts/config_runs.py qv 36 12 2>$O/qv36.err | tee $O/qv36.json
run scripts/config_runs.py qv 33 33 2>$O/qv33.err | tee $O/qv33.json
run bench.py --gpus 8 --steps 2 --warmup 3 2>$O/bench8.err | tee $O/bench8.json | cut -c1-400
