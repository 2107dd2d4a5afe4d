# round 2, call 14 (1 GPU): full GPU suite on the final tree, smoke(), default bench line, reference arm,
# load wall-clock of the 10 M-triangle scene, cost of one rank's band of the strong-scaled 4K frame
T="timeout -k 10"
$T 900 python -m pytest tests -m gpu -q -x --timeout 400 > gpurun_out/r2c14_pytest.log 2>&1; tail -3 gpurun_out/r2c14_pytest.log
$T 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
$T 400 python bench.py > gpurun_out/r2c14_bench_default.json 2> gpurun_out/r2c14_bench_default.err; tail -2 gpurun_out/r2c14_bench_default.err; cut -c1-300 gpurun_out/r2c14_bench_default.json
$T 300 python tools/load_time.py 2236 gpurun_out/r2_load_time_10M.json 2>&1 | tail -1
for n in 8 4 2; do $T 120 python tools/band_probe.py 3840 2160 8 $n 2>&1 | tail -1 | tee -a gpurun_out/r2_band_probe.jsonl; done
$T 120 python tools/band_probe.py 1920 1080 3 1 2>&1 | tail -1 | tee -a gpurun_out/r2_band_probe.jsonl
