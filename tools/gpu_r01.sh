set -x
P=nvidia-gpu-operator-k8s-cluster_b200
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err; tail -3 gpurun_out/bench_r01.err; cat gpurun_out/bench_r01.json
$P/bin/vectorAdd; echo "vectorAdd rc=$?"
mkdir -p /tmp/b200v_val; COMPONENT=cuda OUTPUT_DIR=/tmp/b200v_val $P/bin/nvidia-validator; echo "validator rc=$?"; ls /tmp/b200v_val
COMPONENT=driver OUTPUT_DIR=/tmp/b200v_val $P/bin/nvidia-validator | tail -3; echo "driver rc=$?"; ls /tmp/b200v_val
B200V_MODE=stream $P/bin/nvidia-validator > gpurun_out/stream_r01.jsonl 2> gpurun_out/stream_r01.err; echo "stream rc=$?"; tail -4 gpurun_out/stream_r01.jsonl; tail -3 gpurun_out/stream_r01.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ew_vec_kernel -s 5 -c 2 -f -o gpurun_out/prof_r01 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
