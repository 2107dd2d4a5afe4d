# round 2, call 8: BVH of small scenes in shared memory (fused path kernels) -- parity first, then A/B
T="timeout -k 10"
$T 900 python -m pytest tests -m gpu -q -x --timeout 400 > gpurun_out/r2c8_pytest.log 2>&1; tail -3 gpurun_out/r2c8_pytest.log
rm -f gpurun_out/r2c8_probe.jsonl
V=/root/repo/cuda-pathtracer_b200/_variants
for rep in 1 2; do
for sc in "indoor_tex 64 1920 1080 3" "indoor_tex 16 3840 2160 8" "color_sample_plain 16 3840 2160 4" "indoor_tex 64 640 360 4"; do
  PTC_TAG=global PTC_SMEM_SCENE=0 $T 150 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c8_probe.jsonl
  PTC_TAG=smem   PTC_SMEM_SCENE=1 $T 150 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c8_probe.jsonl
  PTC_TAG=smem3cta PTC_LIBRARY=$V/sms3.so $T 150 python tools/probe.py $sc 2>&1 | tail -1 >> gpurun_out/r2c8_probe.jsonl
done
done
cut -c1-330 gpurun_out/r2c8_probe.jsonl
$T 500 ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend' --launch-skip 12 --launch-count 6 \
   -o gpurun_out/prof_r2_frame_sms -f python tools/frames_only.py indoor_tex 3 1920 1080 3 > gpurun_out/ncu_full_sms.log 2>&1; tail -1 gpurun_out/ncu_full_sms.log | cut -c1-200
$T 400 python bench.py --workload cfg3 --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/r2c8_bench_cfg3_10M.json 2> gpurun_out/r2c8_bench_cfg3.err; tail -2 gpurun_out/r2c8_bench_cfg3.err
