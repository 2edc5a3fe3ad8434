# Sweep of --root-share (rank 0's shard as a fraction of an equal one) on one box.  Usage: gpu_r2_share.sh N share [share ...]
set -x
mkdir -p gpurun_out
N=$1; shift
for S in "$@"; do
for SC in weak strong; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-extra --scaling $SC --root-share $S > gpurun_out/share_${N}_${S}_$SC.log 2>&1
grep '^{' gpurun_out/share_${N}_${S}_$SC.log | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('N=$N share=$S $SC', round(d['ms_per_step'],4), '%.3e'%d['value'], 'collect', round(d['exchange']['collect_ms'],3), 'outside', round(d['exchange']['ms'],3), 'k1', round(d['roofline']['kernel_ms'],3), 'k3', round(d['roofline']['place_kernels']['ms'],3))"
done
done
