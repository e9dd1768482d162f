#!/bin/bash
# Hybrid (switch-side + peer-store) all-reduce on N GPUs: correctness, then share x CTA sweep.
N=${1:-4}
OUT=gpurun_out/r2_hybrid_n$N
V=nvidia-gpu-operator-k8s-cluster_b200/bin/nvidia-validator
export OUTPUT_DIR=/tmp/b200v-out; mkdir -p $OUTPUT_DIR gpurun_out
timeout 400 python -m pytest tests/test_gpu_nvls.py tests/test_gpu_allreduce_comm.py -x -q -k "hybrid or ragged" > $OUT.tests.log 2>&1
echo "tests rc=$?" >> $OUT.tests.log
: > $OUT.sweep.jsonl
run() {  # pct mm_ctas ctas log2
  local line
  line=$(B200V_MODE=nvls B200V_GPUS=$N B200V_NVLS_ALGO=hybrid B200V_NVLS_HYBRID_PCT=$1 B200V_NVLS_HYBRID_MM_CTAS=$2 B200V_NVLS_CTAS=$3 \
         B200V_MIN_LOG2=$4 B200V_MAX_LOG2=$4 timeout 60 $V 2>> $OUT.err | tail -1)
  echo "{\"pct\":$1,\"mm_ctas\":$2,\"ctas\":$3,\"row\":${line:-null}}" >> $OUT.sweep.jsonl
}
for pct in 50 65 75 83 90; do run $pct 16 148 30; done
for pct in 65 83; do run $pct 32 148 30; run $pct 16 100 30; run $pct 8 148 30; done
for pct in 65 83; do run $pct 16 148 26; run $pct 16 148 28; done
tail -3 $OUT.tests.log; cat $OUT.sweep.jsonl | cut -c1-230
