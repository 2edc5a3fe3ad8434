# Weak-scaled tick against the shape of the fused push (tiles per CTA).  Usage: gpu_r2_pushvar.sh N variant [variant ...]
set -x
mkdir -p gpurun_out
N=$1; shift
for V in "$@"; do
TT_PUSH_VARIANT=$V timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e --no-extra > gpurun_out/pushvar_${N}_$V.log 2>&1
grep '^{' gpurun_out/pushvar_${N}_$V.log | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('N=$N variant=$V', round(d['ms_per_step'],4), '%.3e'%d['value'], 'collect', round(d['exchange']['collect_ms'],3), 'outside', round(d['exchange']['ms'],3), 'k1', round(d['roofline']['kernel_ms'],3), 'k3', round(d['roofline']['place_kernels']['ms'],3))"
done
