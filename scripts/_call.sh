set -x
O=gpurun_out/r1e; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/gpu_tests.log 2>&1; tail -5 $O/gpu_tests.log
timeout 300 python scripts/profile_run.py su2 > $O/su2.log 2>&1; grep "\[1\]" $O/su2.log
timeout 300 python scripts/profile_run.py stream > $O/stream.log 2>&1; grep "\[1\]" $O/stream.log
timeout 300 python scripts/profile_run.py sampler > $O/sampler.log 2>&1; grep "\[1\]" $O/sampler.log
python scripts/profile_run.py su2 > $O/su2_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_cluster_pass -s 2 -c 3 -o $O/prof_cluster python scripts/profile_run.py su2 > $O/prof_cluster.log 2>&1
python scripts/profile_run.py stream > $O/stream_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_cluster_pass -s 1 -c 2 -o $O/prof_cluster_qv python scripts/profile_run.py stream > $O/prof_cluster_qv.log 2>&1
