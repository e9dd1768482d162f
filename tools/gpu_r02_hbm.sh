#!/bin/bash
# Round-2 HBM work on one GPU:  tools/gpu_r02_hbm.sh
#   1. array-placement (address phase) experiment and 1R:1W burst shapes (b200v-hbm-exp)
#   2. the tile sweep at 2^24 and 2^26 elements (are the small HBM rows a shape problem?)
#   3. ncu: DRAM counters of one triad launch at 2^28 and one at 2^24 (min/avg/max over the DRAM
#      units, elapsed cycles, bytes), after the same command ran clean without ncu
B=nvidia-gpu-operator-k8s-cluster_b200/bin
OUT=gpurun_out/r2_hbm
mkdir -p gpurun_out
export OUTPUT_DIR=/tmp/b200v-out; mkdir -p $OUTPUT_DIR
timeout 120 $B/b200v-hbm-exp phase 28 > $OUT.phase.csv 2> $OUT.phase.err
timeout 120 $B/b200v-hbm-exp burst 28 > $OUT.burst.csv 2> $OUT.burst.err
timeout 300 $B/b200v-tune 24 20 tiles > $OUT.tune24.csv 2> $OUT.tune24.top.txt
timeout 300 $B/b200v-tune 26 20 tiles > $OUT.tune26.csv 2> $OUT.tune26.top.txt
M=dram__bytes_read.sum,dram__bytes_write.sum,dram__bytes.sum.per_second,dram__bytes.min.per_second,dram__bytes.max.per_second,dram__bytes.avg.per_second,dram__cycles_active.sum,dram__cycles_active.avg,dram__cycles_active.min,dram__cycles_active.max,dram__cycles_elapsed.avg,dram__cycles_elapsed.avg.per_second,dram__throughput.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_write.sum
for lg in 28 24; do
  export B200V_MODE=stream B200V_MIN_LOG2=$lg B200V_MAX_LOG2=$lg
  timeout 120 $B/nvidia-validator > $OUT.stream$lg.jsonl 2>&1 &&
  timeout 600 ncu --metrics $M --clock-control none -k regex:ew_tile_kernel -s $([ $lg = 28 ] && echo 80 || echo 270) -c 2 \
      --csv --log-file $OUT.ncu_triad_$lg.csv $B/nvidia-validator > $OUT.ncu$lg.log 2>&1
done
unset B200V_MODE B200V_MIN_LOG2 B200V_MAX_LOG2
tail -3 $OUT.phase.csv $OUT.burst.csv; tail -4 $OUT.ncu24.log
