# 1-GPU call: no-alloc tests of the final build, then ONE --set full capture of a whole step's pairconv / wgrad / haloconv
# launches (46 kernels); the raw page is exported to CSV on the box because the report is larger than gpurun copies back
set -x
timeout 100 python -m pytest tests/test_zz_no_alloc.py -q --timeout 80 2>&1 | tail -3
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-e2e"
timeout 60 $CMD > gpurun_out/ncu_plain.log 2>&1 || { tail -5 gpurun_out/ncu_plain.log; exit 1; }
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'pairconv|wgrad|haloconv' \
    --launch-skip 138 --launch-count 46 -f -o gpurun_out/prof_final_r2 $CMD > gpurun_out/ncu_final.log 2>&1
tail -2 gpurun_out/ncu_final.log
ncu -i gpurun_out/prof_final_r2.ncu-rep --page raw --csv > gpurun_out/prof_final_r2_raw.csv 2>/dev/null
ls -la gpurun_out/ | tail -5
SZ=$(stat -c %s gpurun_out/prof_final_r2.ncu-rep); if [ "$SZ" -gt 30000000 ]; then rm gpurun_out/prof_final_r2.ncu-rep; echo "report too large ($SZ), kept the csv only"; fi
du -sh gpurun_out
