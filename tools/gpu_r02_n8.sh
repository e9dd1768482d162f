#!/bin/bash
# Round-2 eight-GPU run (every minute here costs eight):  tools/gpu_r02_n8.sh [N=8]
#   1. host<->device matrix (b200v-linkbw pcie): names what limits the host-buffer path at N >= 2
#   2. the host-buffer path under each host-memory policy (bench.py --e2e-only)
#   3. small-message all-reduce rows: one-shot against two-shot, and the one-shot threshold
#   4. the product's own per-GPU fan-out at 2^28 elements (cfg3)
#   5. bench.py --gpus N (oracle-checked all-reduce parity inside) and the CPU reference arm
N=${1:-8}
OUT=gpurun_out/r2_n$N
B=nvidia-gpu-operator-k8s-cluster_b200/bin
mkdir -p gpurun_out
export OUTPUT_DIR=/tmp/b200v-out; mkdir -p $OUTPUT_DIR
nvidia-smi topo -m > $OUT.topo.txt 2>&1
lscpu | head -40 > $OUT.lscpu.txt 2>&1; numactl -H >> $OUT.lscpu.txt 2>&1; grep -i huge /proc/meminfo >> $OUT.lscpu.txt
cat /sys/kernel/mm/transparent_hugepage/enabled >> $OUT.lscpu.txt 2>&1

timeout 400 $B/b200v-linkbw pcie $N > $OUT.pcie.jsonl 2> $OUT.pcie.err

run_tr() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
: > $OUT.e2e_policy.jsonl
port=29600
for pol in pinned wc thp; do
  port=$((port+1))
  B200V_HOST_POLICY=$pol run_tr $port --e2e-only >> $OUT.e2e_policy.jsonl 2>> $OUT.e2e_policy.err
done

for algo in auto multimem; do
  if [ $algo = auto ]; then unset B200V_NVLS_ALGO; else export B200V_NVLS_ALGO=$algo; fi
  B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=10 B200V_MAX_LOG2=24 timeout 120 $B/nvidia-validator > $OUT.small_$algo.jsonl 2>> $OUT.ar.err
done
unset B200V_NVLS_ALGO
for mx in 65536 1048576; do
  B200V_NVLS_ONESHOT_MAX=$mx B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=14 B200V_MAX_LOG2=21 timeout 120 $B/nvidia-validator > $OUT.oneshot_max_$mx.jsonl 2>> $OUT.ar.err
done
for ctas in 8 24 32; do
  B200V_NVLS_CTAS=$ctas B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=30 B200V_MAX_LOG2=30 timeout 60 $B/nvidia-validator >> $OUT.multimem_ctas.jsonl 2>> $OUT.ar.err
done

B200V_MODE=nvls B200V_GPUS=$N B200V_MIN_LOG2=10 B200V_MAX_LOG2=30 timeout 200 $B/nvidia-validator > $OUT.nvls.jsonl 2>> $OUT.ar.err
timeout 100 $B/b200v-linkbw p2p $N > $OUT.p2p.jsonl 2>> $OUT.ar.err

rm -f $OUTPUT_DIR/*-ready
B200V_ELEMS=268435456 B200V_GPUS=$N B200V_WRITE_PROM=1 COMPONENT=cuda timeout 300 $B/nvidia-validator > $OUT.fanout.log 2>&1
echo "fanout rc=$? files: $(ls $OUTPUT_DIR | tr '\n' ' ')" >> $OUT.fanout.log

timeout 200 python bench.py --impl reference --gpus $N --steps 20 --warmup 5 > $OUT.reference.json 2> $OUT.reference.err
run_tr 29650 --steps 20 --warmup 5 > $OUT.bench.json 2> $OUT.bench.err
echo "bench rc=$?" >> $OUT.bench.err
cat $OUT.e2e_policy.jsonl; grep '"mode":"fanout"' $OUT.fanout.log; tail -c 400 $OUT.bench.err
