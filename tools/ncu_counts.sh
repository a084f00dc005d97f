M=smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 || exit 1
for tl in 0 1; do
HRC_TUNE_TRI_LANES=$tl HRC_TUNE_REFILL_IDLE=10 ncu --metrics $M --clock-control none -k regex:k_trace -s 54 -c 6 --csv --log-file gpurun_out/cnt_t${tl}.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
done
