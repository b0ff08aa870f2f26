# build-time variants of k_fused_pass, timed on the four benchmark streams (nvcc is on the GPU box image)
O=gpurun_out/${1:-r2r}; mkdir -p $O
run() { for m in su2 qv qft tfim; do timeout 300 python scripts/profile_run.py $m 2>&1 | grep "^\[1\]" | cut -c1-70; done; }
echo "== default build"; run | tee $O/default.txt
for def in "$@"; do
  [ "$def" = "$1" ] && continue
  echo "== -D$def"; sed -i "s/NVCC_FLAGS = \[/NVCC_FLAGS = [\"-D$def\", /" qiskit_b200/build.py
  python -m qiskit_b200.build --force > $O/build_$def.log 2>&1 || tail -3 $O/build_$def.log
  run | tee $O/$def.txt
  sed -i "s/NVCC_FLAGS = \[\"-D$def\", /NVCC_FLAGS = [/" qiskit_b200/build.py
done
