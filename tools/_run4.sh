ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_launches.csv -k regex:"conv_tc|conv_mma_small" python tools/bench_small.py > gpurun_out/bench_small.log 2>&1
tail -3 gpurun_out/bench_small.log
