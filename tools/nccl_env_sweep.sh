# NCCL tuning sweep for the single-process probe: busBW at 256 MiB and 1 GiB under different NCCL_* settings.
# Every setting runs under its own 40 s timeout: some NCCL_ALGO / transport overrides hang at small rank counts
# (an unguarded run of this sweep cost a whole 600 s gpurun limit once).
# usage: bash tools/nccl_env_sweep.sh N
N=${1:-2}
V=nvidia-gpu-operator-k8s-cluster_b200/bin/nvidia-validator
run() { echo -n "[$*] "; timeout 40 env "$@" B200V_MODE=nccl B200V_GPUS=$N B200V_MIN_LOG2=28 B200V_MAX_LOG2=30 $V 2>/dev/null | python -c "
import sys, json
rows=[json.loads(l) for l in sys.stdin if l.startswith('{')]
print(' '.join('%d:%.0f' % (r['bytes'].bit_length()-1, r['busbw_gbs']) for r in rows), 'bad', sum(r['mismatches'] for r in rows))"; }
run X=1
run NCCL_MIN_NCHANNELS=32
run NCCL_MIN_NCHANNELS=64 NCCL_MAX_NCHANNELS=64
run NCCL_BUFFSIZE=16777216
run NCCL_BUFFSIZE=67108864 NCCL_MIN_NCHANNELS=32
run NCCL_NTHREADS=512
run NCCL_PROTO=Simple NCCL_ALGO=Ring
run NCCL_ALGO=Tree
run NCCL_ALGO=NVLS
run NCCL_P2P_USE_CUDA_MEMCPY=1
run NCCL_NVLS_ENABLE=0
run NCCL_CUMEM_ENABLE=1
