# A/B of one engine env switch: bash scripts/ab_env.sh <outdir> <ENVNAME> [modes...]
O=gpurun_out/$1; E=$2; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/tests.log 2>&1; echo "tests exit $?"; tail -3 $O/tests.log
for v in 0 1; do
  echo "== $E=$v"
  env $E=$v timeout 300 python scripts/profile_run.py su2 2>&1 | cut -c1-150 | tee $O/su2_$v.txt
  env $E=$v QSV_TRACE=1 timeout 300 python scripts/trace_run.py su2 28 > $O/trace_su2_$v.txt 2>&1
  env $E=$v timeout 300 python scripts/profile_run.py sampler 2>&1 | cut -c1-150 | tee $O/sampler_$v.txt
done
