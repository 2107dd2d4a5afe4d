# parity tests, then one --set full capture of the three stage kernels of one bench frame
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu.log 2>&1
grep -n "^E  \|^____\|passed\|failed" gpurun_out/pytest_gpu.log | grep -v "where\|+ " | head -40
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_path|k_extend' --launch-skip 30 --launch-count 3 \
    -o gpurun_out/prof_r1_v5_frame -f python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
