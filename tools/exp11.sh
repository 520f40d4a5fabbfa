set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -40 > gpurun_out/gputest.log; tail -5 gpurun_out/gputest.log
python bench.py --steps 100 --warmup 5 2>gpurun_out/bench.err | tail -1 > gpurun_out/bench_r02_a.json; cat gpurun_out/bench_r02_a.json
python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 > gpurun_out/bench_r02_a_reference.json; cat gpurun_out/bench_r02_a_reference.json
