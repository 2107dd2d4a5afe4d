# full ncu capture (with source counters) of the stage kernels of one headline frame, one pipeline
PTC_DUAL=0 python tools/frame_probe.py indoor_tex 6 > gpurun_out/plain_probe.log 2>&1 && \
PTC_DUAL=0 ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend|k_wave' --launch-skip 15 --launch-count 3 \
    -o gpurun_out/prof_src -f python tools/frame_probe.py indoor_tex 6 > gpurun_out/ncu_src.log 2>&1
tail -2 gpurun_out/ncu_src.log | cut -c1-200
