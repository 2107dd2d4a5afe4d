# round measurements: GPU tests, bench lines of all single-GPU configs, launch list and full capture of the frame
bash tools/gpu_measure.sh
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend|k_wave' --launch-skip 60 --launch-count 6 \
    -o gpurun_out/prof_r1_v8_frame -f python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log | cut -c1-200
