set -x
O=gpurun_out/r1f; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s > $O/dist_tests.log 2>&1; tail -40 $O/dist_tests.log
