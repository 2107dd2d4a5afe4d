# round 2, call 24: load wall-clock of the 10 M-triangle scene with the restructured host builder (16 host threads)
PTC_BUILD_TIMING=1 timeout -k 5 60 python tools/load_time.py 2236 gpurun_out/r2_load_time_10M_new_builder.json 2>&1 | grep -E "build_bvh|faces" | tail -8
