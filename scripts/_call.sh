set -x
mkdir -p gpurun_out/r1
O=gpurun_out/r1
python scripts/profile_run.py su2 > $O/su2_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/su2_launches.csv python scripts/profile_run.py su2 > $O/su2_ncu.log 2>&1
python scripts/profile_run.py stream > $O/stream_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/stream_launches.csv python scripts/profile_run.py stream > $O/stream_ncu.log 2>&1
python scripts/profile_run.py sampler > $O/sampler_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/sampler_launches.csv python scripts/profile_run.py sampler > $O/sampler_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile_pass -s 2 -c 3 -o $O/prof_tile python scripts/profile_run.py su2 > $O/prof_tile.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_direct2 -s 1 -c 2 -o $O/prof_direct2 python scripts/profile_run.py stream > $O/prof_direct2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_expval -c 3 -o $O/prof_expval python scripts/profile_run.py su2 > $O/prof_expval.log 2>&1
python scripts/tune_gate.py 30 > $O/tune30.log 2>&1
cat $O/su2_plain.log $O/stream_plain.log $O/sampler_plain.log
ls -la $O
