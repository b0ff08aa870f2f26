# Round 2, final 8-GPU check: world-8 sharded parity test, the driver's N = 8 bench command (full), N = 4 sharded block.
O=gpurun_out/${1:-r2j8}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s -k "8" > $O/dist8.log 2>&1; echo "dist world-8 exit $?"; grep -c "amp err" $O/dist8.log; grep "named10\|qv20" $O/dist8.log | head -8; tail -3 $O/dist8.log | cut -c1-300
grep -B2 -A12 "Error\|assert" $O/dist8.log | grep -v "^E   *\(time\|host\|exitcode\|error_file\|traceback\|rank\)" | head -40
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 8 --steps 2 --warmup 3 > $O/bench8.json 2> $O/bench8.err; echo "bench8 exit $?"
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29702 bench.py --gpus 4 --steps 1 --warmup 1 --skip-extras > $O/bench4.json 2> $O/bench4.err; echo "bench4 exit $?"
python - <<'PY'
import json
for f in ("bench8", "bench4"):
    try:
        d = json.loads([l for l in open(f"/root/repo/gpurun_out/r2j8/{f}.json") if l.startswith("{")][-1])
        s = d["sharded"]
        print(f, "value", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), "parity", s["parity"].get("ok"), s["parity"].get("max_abs_amp_err"),
              "qv_strong ms", round(s["qv_strong"]["gate_ms"], 1), "swaps", s["qv_strong"]["swap_count"], "fused", s["qv_strong"]["fused_swaps"],
              "nvlink", round(s["qv_strong"].get("nvlink_gbs_per_direction", 0), 1))
        if "qv_big" in s:
            print("  qv_big ms", round(s["qv_big"]["gate_ms"], 1), "eff", round(s["qv_big"]["parallel_efficiency_time"], 3), round(s["qv_big"]["parallel_efficiency_gbs"], 3),
                  "tfim ms", round(s["tfim_big"]["gate_ms"], 1), "energy", s["tfim_big"]["energy"])
    except Exception as e:
        print(f, "no json", e)
PY
grep -c "NCCL INFO" $O/bench8.err; grep "nranks" $O/bench8.err | head -2 | cut -c1-200
