# ncu capture of the library's other kernels (a few launches each, after their first) with the metrics the
# roofline table needs: DRAM bytes, duration, issue / warp activity, L1 data-pipe load, occupancy limits.
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__occupancy_limit_registers,launch__occupancy_limit_shared_mem,launch__grid_size,launch__block_size
python bench_kernels.py --out gpurun_out/kernels_r02.json > gpurun_out/kernels_r02.txt 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'parse_samples|tokd_|series_partial|group_fold|group_step|hist_kernel|pick_kernel|tail_kernel|init_kernel|finish_kernel|ragged_fast|ragged_window|bucket_kernel|slab_|window_fast' -c 400 --csv --log-file gpurun_out/ncu_kernels_r02.csv python bench_kernels.py > /dev/null 2>&1
tail -30 gpurun_out/kernels_r02.txt
