O=gpurun_out/${1:-r2l}; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 2 --steps 1 --warmup 1 --skip-extras --sharded-qubits 30 > $O/bench2.out 2> $O/bench2.err; echo "bench2 exit $?"
grep -c "" $O/bench2.out; grep -v "^{" $O/bench2.out | cut -c1-200; tail -n 1 $O/bench2.out | cut -c1-120; grep "NCCL" $O/bench2.err | cut -c1-200
