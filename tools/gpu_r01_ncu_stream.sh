# Round-1 ncu evidence for the four STREAM ops at 2^28 elements (1 GiB per array): one full-set
# capture of one HBM-bound launch per op.  1 GPU.  The plain run of the same command comes first.
# The sweep launches the elementwise kernel 25 times per op (5 warm-up + 20 timed) in the order
# copy, scale, add, triad, so launch 5 / 30 / 55 / 80 of ew_tile_kernel is a warm launch of each.
mkdir -p gpurun_out
V=nvidia-gpu-operator-k8s-cluster_b200/bin/nvidia-validator
export B200V_MODE=stream B200V_MIN_LOG2=28 B200V_MAX_LOG2=28
timeout 120 $V > gpurun_out/stream28_plain.jsonl 2> gpurun_out/stream28_plain.err || exit 1
cut -c1-200 gpurun_out/stream28_plain.jsonl
for pair in copy:5 scale:30 add:55 triad:80; do
  op=${pair%%:*}; skip=${pair##*:}
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:ew_tile_kernel -s $skip -c 1 -f \
      -o gpurun_out/prof_r01_stream_$op $V > gpurun_out/ncu_stream_$op.log 2>&1
  echo "ncu $op rc=$? $(grep -c '==PROF== Profiling' gpurun_out/ncu_stream_$op.log) kernel(s)"
done
