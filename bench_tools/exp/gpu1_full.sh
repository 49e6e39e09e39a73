# 1-GPU call: whole GPU test suite, launch list of one step, --set full captures of the top kernels (small reports only:
# gpurun copies back at most 64 MiB)
set -x
timeout 600 python -m pytest tests -q -m gpu --timeout 300 > gpurun_out/gputests.log 2>&1; tail -12 gpurun_out/gputests.log
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-e2e"
timeout 120 $CMD > gpurun_out/ncu_plain.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:'conv|bn_|head|stem|push|pull|wait|bump|gate' --launch-skip 372 --launch-count 128 --csv --log-file gpurun_out/launches_v6.csv $CMD > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'pairconv|wgrad_halo|haloconv|head_fused|tapconv' \
    --launch-skip 150 --launch-count 14 -f -o gpurun_out/prof_top_r2 $CMD > gpurun_out/ncu_top.log 2>&1
tail -2 gpurun_out/ncu_top.log
ncu -i gpurun_out/prof_top_r2.ncu-rep --page raw --csv > gpurun_out/prof_top_r2_raw.csv 2>/dev/null
ls -la gpurun_out/; SZ=$(stat -c %s gpurun_out/prof_top_r2.ncu-rep); if [ "$SZ" -gt 45000000 ]; then rm gpurun_out/prof_top_r2.ncu-rep; echo "report too large, kept the csv only"; fi
