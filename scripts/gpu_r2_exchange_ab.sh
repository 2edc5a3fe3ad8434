# Several exchange configurations in one torchrun launch.  Usage: gpu_r2_exchange_ab.sh N spec [spec ...]  (spec = form:share:scaling)
set -x
mkdir -p gpurun_out
N=$1; shift
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 scripts/exchange_ab.py "$@" > gpurun_out/early_ab_$N.log 2>&1
grep '^{' gpurun_out/early_ab_$N.log | cut -c1-620
tail -3 gpurun_out/early_ab_$N.log | cut -c1-300
