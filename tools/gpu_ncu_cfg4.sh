python bench.py --steps 2 --warmup 3 --scene synth2236 --bounces 4 --no-cpu-baseline > gpurun_out/plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend' --launch-skip 28 --launch-count 14 \
    -o gpurun_out/prof_r1_v8_cfg4 -f python bench.py --steps 2 --warmup 3 --scene synth2236 --bounces 4 --no-cpu-baseline > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu4.log | cut -c1-300
