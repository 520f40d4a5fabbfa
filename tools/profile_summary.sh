# ncu: launch list of two bench steps, and the kernels of the one-sweep summary (one rank and the
# several-rank form on a one-rank group) with duration, DRAM bytes, issue / warp activity, occupancy.
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__grid_size,launch__block_size
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-config3 > gpurun_out/b2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r02.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-config3 > /dev/null 2>&1
python tools/run_summary_once.py > /dev/null 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'bs_|group_fold|window_fast' -s 6 -c 5 --csv --log-file gpurun_out/summary_r02_ncu.csv python tools/run_summary_once.py > /dev/null 2>&1
python tools/run_summary_ranks_once.py > /dev/null 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'bs_|group_fold' -s 26 -c 13 --csv --log-file gpurun_out/summary_ranks_r02_ncu.csv python tools/run_summary_ranks_once.py > /dev/null 2>&1
ls -la gpurun_out/*.csv
