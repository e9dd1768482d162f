# Multi-GPU round-1 run: usage  bash tools/gpu_r01_multi.sh N   (under gpurun --gpus N)
# Writes everything under gpurun_out/: GPU test log, the validator fan-out, the NCCL probe at
# 2/4/../N ranks and bench.py at 1/2/4/../N GPUs plus the CPU reference arm.
N=${1:-2}
P=nvidia-gpu-operator-k8s-cluster_b200
mkdir -p gpurun_out
nvidia-smi topo -m | head -12 > gpurun_out/topo_n$N.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4

rm -rf /tmp/b200v_val; mkdir -p /tmp/b200v_val
COMPONENT=cuda OUTPUT_DIR=/tmp/b200v_val B200V_ELEMS=268435456 $P/bin/nvidia-validator > gpurun_out/fanout_n$N.log 2>&1
echo "fanout rc=$? files: $(ls /tmp/b200v_val | tr '\n' ' ')" | tee -a gpurun_out/fanout_n$N.log

for g in 2 4 8; do
  if [ $g -le $N ]; then
    B200V_MODE=nccl B200V_GPUS=$g $P/bin/nvidia-validator > gpurun_out/nccl_r01_n$g.jsonl 2> gpurun_out/nccl_r01_n$g.err
    echo "nccl ranks=$g rc=$? $(tail -1 gpurun_out/nccl_r01_n$g.jsonl | cut -c1-190)"
    B200V_NCCL_PLAIN_BUFFERS=1 B200V_MODE=nccl B200V_GPUS=$g B200V_MIN_LOG2=28 $P/bin/nvidia-validator 2>&1 | tail -1 | cut -c1-190 | sed 's/^/   plain cudaMalloc buffers: /'
  fi
done

for k in 1 2 4 8; do
  if [ $k -le $N ]; then
    if [ $k -eq 1 ]; then python bench.py --gpus 1 --steps 100 --warmup 5 > gpurun_out/scale_r01_n1.json 2> gpurun_out/scale_r01_n1.err
    else python -m torch.distributed.run --nnodes=1 --nproc-per-node $k --master-addr 127.0.0.1 --master-port 2951$k bench.py --gpus $k --steps 100 --warmup 5 > gpurun_out/scale_r01_n$k.json 2> gpurun_out/scale_r01_n$k.err; fi
    echo "bench n=$k rc=$?"; tail -2 gpurun_out/scale_r01_n$k.err | cut -c1-300
    python - <<EOF
import json
d = json.loads(open("gpurun_out/scale_r01_n$k.json").read().strip().splitlines()[-1])
ar = d.get("allreduce") or {}
print("  n=%d value=%.1f per_gpu=%.1f triad=%.1f e2e=%.2f busbw@1GiB=%s clocks=%s" % (
    d["n_gpus"], d["value"], d["per_gpu_gbs"], d["triad_gbs"], d["e2e"]["value"], ar.get("busbw_gbs_at_1GiB"), d["clocks"]))
EOF
  fi
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520 bench.py --impl reference --gpus $N --steps 10 --warmup 2 2>/dev/null | tail -1 > gpurun_out/reference_r01_n$N.json
cut -c1-330 gpurun_out/reference_r01_n$N.json
