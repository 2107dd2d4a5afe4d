# round 2, call 19: pipelines per frame by size, refill threshold of the extend kernel (both on the adopted schedule)
T="timeout -k 10"
$T 400 python tools/sweep_dual.py 2>&1 | grep '^{' | tee gpurun_out/r2_sweep_dual.jsonl
$T 120 python tools/sweep_refill.py synth2236 4 1920 1080 2>&1 | grep refill | tee gpurun_out/r2_sweep_refill.txt
$T 120 python tools/sweep_refill.py synth300 4 1920 1080 2>&1 | grep refill | tee -a gpurun_out/r2_sweep_refill.txt
