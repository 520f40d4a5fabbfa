# ncu metrics pass over the kernels changed late in round 2: device tokenizer, K1 decode, ragged kernel
# (each command first runs to completion without the profiler).
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__occupancy_limit_registers,launch__occupancy_limit_shared_mem,launch__grid_size,launch__block_size
python tools/run_tokenize_once.py > gpurun_out/tok_once.txt 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'tokd_' -s 9 -c 3 --csv --log-file gpurun_out/ncu_tokd_r02.csv python tools/run_tokenize_once.py > /dev/null 2>&1
python tools/run_parse_once.py > gpurun_out/parse_once.txt 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'parse_samples' -s 3 -c 1 --csv --log-file gpurun_out/ncu_parse_r02.csv python tools/run_parse_once.py > /dev/null 2>&1
python tools/time_ragged.py 100000 > gpurun_out/ragged_once.txt 2>&1 &&
ncu --metrics $M --clock-control none -k regex:'ragged_fast' -s 2 -c 12 --csv --log-file gpurun_out/ncu_ragged_r02.csv python tools/time_ragged.py 100000 > /dev/null 2>&1
cat gpurun_out/tok_once.txt gpurun_out/parse_once.txt gpurun_out/ragged_once.txt
