#!/bin/bash
# Round-2 closing run on one GPU: the whole GPU suite, smoke(), the bench line, then the ncu
# launch list and one --set full capture of the headline kernel from a short bench run (each ncu
# pass only after the same command exited 0 without it).
mkdir -p gpurun_out
OUT=gpurun_out/r2_final
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -15 > $OUT.gputest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $OUT.smoke.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 5 > $OUT.bench_n1.json 2> $OUT.bench_n1.err
SHORT="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-stream-sweep"
timeout 300 $SHORT > $OUT.short.json 2> $OUT.short.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT.launches.csv $SHORT > $OUT.ncu1.log 2>&1
timeout 300 $SHORT > $OUT.short2.json 2>> $OUT.short.err &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ew_tile_kernel -s 5 -c 2 -o $OUT.vectoradd_tile $SHORT > $OUT.ncu2.log 2>&1
tail -4 $OUT.gputest.log; cat $OUT.smoke.log; tail -c 300 $OUT.bench_n1.err; tail -2 $OUT.ncu1.log $OUT.ncu2.log
