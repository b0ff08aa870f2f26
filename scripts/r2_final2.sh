# Round 2, final 2-GPU check of both bench arms under torchrun (NCCL INFO lines on rank 0, reference arm with all host threads).
O=gpurun_out/${1:-r2k}; mkdir -p $O
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 1 --warmup 1 --skip-extras > $O/bench2.out 2> $O/bench2.err; echo "bench2 exit $?"
grep -c "NCCL INFO" $O/bench2.out $O/bench2.err; grep -h "nranks" $O/bench2.out $O/bench2.err | head -3 | cut -c1-220; tail -n 1 $O/bench2.out | cut -c1-300
python - <<'PY'
import json
lines = open('/root/repo/gpurun_out/r2k/bench2.out').read().splitlines()
d = json.loads(lines[-1]); print("last stdout line is the JSON line:", d["metric"], d["n_gpus"], d["sharded"]["parity"]["ok"], d["sharded"]["qv_strong"]["gate_ms"])
PY
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29722 bench.py --gpus 2 --steps 3 --warmup 1 --impl reference > $O/ref2.out 2> $O/ref2.err; echo "ref2 exit $?"; tail -n 1 $O/ref2.out | cut -c1-400
python - <<'PY'
import json
d = json.loads(open('/root/repo/gpurun_out/r2k/ref2.out').read().splitlines()[-1])
print("threads", d["config"]["threads"], "value", d["value"], "ms_per_step", d["ms_per_step"], "config0 evals/s", d["config0"]["evals_per_s"])
print("model_check", d["model_check"])
PY
