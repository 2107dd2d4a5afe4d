python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu.log 2>&1
grep -n "^E  \|^____\|passed\|failed" gpurun_out/pytest_gpu.log | grep -v "where\|+ " | head -40
python tools/frame_probe.py indoor_tex 20; python tools/sweep.py
python bench.py --steps 32 --warmup 4 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; tail -3 gpurun_out/bench1.err; cat gpurun_out/bench1.json
