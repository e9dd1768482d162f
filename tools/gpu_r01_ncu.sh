# Round-1 ncu evidence for the shipped default kernel (ew_tile_kernel).  1 GPU.
#   launch list:  every launch of a short bench run with its device time
#   full capture: two full-size vectorAdd launches with --set full
# Each ncu run follows a plain run of the same command line that exited 0 (B200_PROFILING.md).
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_a.log 2> gpurun_out/plain_a.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_r01b.csv $CMD > gpurun_out/ncu_list_b.log 2>&1
echo "ncu list rc=$?"
$CMD > gpurun_out/plain_b.log 2> gpurun_out/plain_b.err &&
ncu --set full --clock-control none --import-source on -k regex:ew_tile_kernel -s 5 -c 2 -f -o gpurun_out/prof_r01b $CMD > gpurun_out/ncu_full_b.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_full_b.log
python bench.py > gpurun_out/bench_r01b.json 2> gpurun_out/bench_r01b.err; echo "bench rc=$?"; wc -l gpurun_out/bench_r01b.json; cut -c1-400 gpurun_out/bench_r01b.json
