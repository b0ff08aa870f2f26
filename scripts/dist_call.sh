# usage: bash scripts/dist_call.sh <world> <outdir> <mode> <n> <depth> [extra env...]
W=$1; O=$2; shift 2
mkdir -p $O
python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29611 scripts/config_runs.py "$@"
