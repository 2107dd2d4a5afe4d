python - <<'PY'
import sys; sys.path.insert(0, '.')
import bench
P = bench.pkg(); c = P.Core(0)
for ws in (16 << 10, 32 << 10, 64 << 10, 128 << 10, 1 << 20, 32 << 20, 64 << 20, 256 << 20, 4 << 30):
    print("working set %10d B: %8.0f GB/s" % (ws, c.read_bandwidth(ws, 16 << 30)))
PY
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend' --launch-skip 30 --launch-count 3 \
    -o gpurun_out/prof_r1_v6_frame -f python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log | cut -c1-200
