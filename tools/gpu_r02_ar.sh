#!/bin/bash
# Round-2 all-reduce work on N GPUs (N = 2 or 4):  tools/gpu_r02_ar.sh N
#   1. multi-GPU GPU tests (hand-written probes, caller-owned-buffer comms vs the oracle)
#   2. peer-copy link rate (b200v-linkbw p2p)
#   3. kernel x CTA x unroll sweep of the hand-written all-reduce at 1 MiB / 64 MiB / 1 GiB
#   4. small-message latency rows, one-shot vs two-shot
#   5. NVLink TX/RX counters around a run of 1 GiB all-reduces (traffic evidence)
#   6. bench.py --gpus N
# Everything is bounded by its own timeout; results land in gpurun_out/r2_ar_n$N.*
N=${1:-2}
OUT=gpurun_out/r2_ar_n$N
V=nvidia-gpu-operator-k8s-cluster_b200/bin/nvidia-validator
L=nvidia-gpu-operator-k8s-cluster_b200/bin/b200v-linkbw
mkdir -p gpurun_out
export OUTPUT_DIR=/tmp/b200v-out; mkdir -p $OUTPUT_DIR
nvidia-smi topo -m > $OUT.topo.txt 2>&1

timeout 900 python -m pytest tests/test_gpu_nvls.py tests/test_gpu_allreduce_comm.py tests/test_zz_no_alloc.py \
    "tests/test_gpu_parity.py::test_allreduce_probe_all_devices" -x -q > $OUT.tests.log 2>&1
echo "tests rc=$?" >> $OUT.tests.log

timeout 120 $L p2p $N > $OUT.p2p.jsonl 2> $OUT.p2p.err

: > $OUT.sweep.jsonl
run() {  # algo ctas unroll log2
  local line
  line=$(B200V_MODE=nvls B200V_GPUS=$N B200V_NVLS_ALGO=$1 B200V_NVLS_CTAS=$2 B200V_NVLS_UNROLL=$3 \
         B200V_MIN_LOG2=$4 B200V_MAX_LOG2=$4 timeout 60 $V 2>> $OUT.sweep.err | tail -1)
  echo "{\"algo\":\"$1\",\"ctas\":$2,\"unroll\":$3,\"row\":${line:-null}}" >> $OUT.sweep.jsonl
}
for algo in push p2p multimem; do
  for ctas in 8 16 32 64 128 148; do
    for u in 2 4; do run $algo $ctas $u 30; done
  done
  for ctas in 16 32 64 128; do run $algo $ctas 4 26; done
  for ctas in 8 32; do run $algo $ctas 4 20; done
done
for ctas in 32 64; do run push $ctas 8 30; run push $ctas 1 30; done

# small messages: the default (one-shot up to its threshold) against forced two-shot kernels
for algo in auto multimem push; do
  if [ $algo = auto ]; then unset B200V_NVLS_ALGO; else export B200V_NVLS_ALGO=$algo; fi
  B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=10 B200V_MAX_LOG2=24 timeout 120 $V > $OUT.small_$algo.jsonl 2>> $OUT.sweep.err
done
unset B200V_NVLS_ALGO
for mx in 131072 1048576; do
  B200V_NVLS_ONESHOT_MAX=$mx B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=14 B200V_MAX_LOG2=21 timeout 120 $V > $OUT.oneshot_max_$mx.jsonl 2>> $OUT.sweep.err
done
B200V_MODE=nccl B200V_GPUS=$N B200V_MIN_LOG2=10 B200V_MAX_LOG2=30 timeout 300 $V > $OUT.nccl.jsonl 2>> $OUT.sweep.err
B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=10 B200V_MAX_LOG2=30 timeout 300 $V > $OUT.nvls.jsonl 2>> $OUT.sweep.err

# NVLink byte counters around 20+ all-reduces of 1 GiB (per-link data TX/RX in KiB)
nvidia-smi nvlink -gt d -i 0 > $OUT.nvlink_before.txt 2>&1
B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=30 B200V_MAX_LOG2=30 timeout 120 $V > $OUT.nvlink_run.jsonl 2>> $OUT.sweep.err
nvidia-smi nvlink -gt d -i 0 > $OUT.nvlink_after.txt 2>&1

timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 > $OUT.bench.json 2> $OUT.bench.err
echo "bench rc=$?" >> $OUT.bench.err
tail -3 $OUT.tests.log; wc -l $OUT.sweep.jsonl; tail -c 300 $OUT.bench.err
