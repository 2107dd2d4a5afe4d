# round 2, call 22: load wall-clock of the 10 M-triangle scene with the faster host build; the large-mesh parity tests on it
T="timeout -k 10"
PTC_BUILD_TIMING=1 $T 200 python tools/load_time.py 2236 gpurun_out/r2_load_time_10M.json 2>&1 | grep -E "build_bvh|faces" | tail -12
$T 300 python -m pytest tests -m gpu -q -x --timeout 250 -k "config4 or large_mesh or traversal_counters or hits_bit_identical" 2>&1 | tail -2
