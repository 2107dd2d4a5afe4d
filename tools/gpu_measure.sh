set -x
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; tail -1 gpurun_out/pytest_gpu.log
python bench.py --steps 64 --warmup 8 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
python bench.py --impl reference --steps 8 --warmup 2 > gpurun_out/bench_ref_n1.json 2>> gpurun_out/bench_n1.err
python bench.py --steps 16 --warmup 4 --scene indoor_plain > gpurun_out/bench_plain.json 2>> gpurun_out/bench_n1.err
python bench.py --steps 8 --warmup 3 --width 640 --height 360 --bounces 4 --no-cpu-baseline > gpurun_out/bench_cfg1_640x360_b4.json 2>> gpurun_out/bench_n1.err
python bench.py --steps 16 --warmup 3 --bounces 4 --no-cpu-baseline > gpurun_out/bench_cfg2_1080p_b4.json 2>> gpurun_out/bench_n1.err
python bench.py --steps 16 --warmup 3 --width 3840 --height 2160 --bounces 8 --no-cpu-baseline > gpurun_out/bench_cfg3_4k_b8_1gpu.json 2>> gpurun_out/bench_n1.err
python bench.py --steps 16 --warmup 3 --scene synth2236 --bounces 4 --no-cpu-baseline > gpurun_out/bench_cfg4_10M.json 2>> gpurun_out/bench_n1.err
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/bench_n1.err
