# round 2, call 10 (2 GPUs): host-image e2e path at N = 2, new tests
T="timeout -k 10"
$T 600 python -m pytest tests -m gpu -q -x --timeout 400 -k "strips_scattered or group or full_image or shared_memory" > gpurun_out/r2c10_pytest.log 2>&1; tail -3 gpurun_out/r2c10_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
$T 400 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2c10_bench_n2.log 2> gpurun_out/r2c10_bench_n2.err; tail -3 gpurun_out/r2c10_bench_n2.err; grep '^{' gpurun_out/r2c10_bench_n2.log > gpurun_out/r2_scale_n2.json; cut -c1-300 gpurun_out/r2_scale_n2.json
$T 400 $TR bench.py --gpus 2 --workload cfg2 --steps 16 --warmup 3 > gpurun_out/r2c10_cfg2_n2.log 2> gpurun_out/r2c10_cfg2_n2.err; tail -3 gpurun_out/r2c10_cfg2_n2.err; grep '^{' gpurun_out/r2c10_cfg2_n2.log > gpurun_out/r2_cfg2_n2.json; cut -c1-300 gpurun_out/r2_cfg2_n2.json
python - <<'PY'
import json
for f in ("r2_scale_n2", "r2_cfg2_n2"):
    d = json.load(open(f"gpurun_out/{f}.json"))
    print(f, d["ms_per_step"], d["value"], d["e2e"], d["bit_identity"])
PY
ls /dev/shm | head
