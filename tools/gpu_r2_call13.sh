# round 2, call 13 (8 GPUs): DMA-copy assembly at N = 8 and N = 4 (headline weak scaling, configs[2] strong scaling + bit identity)
T="timeout -k 10"
run() { N=$1; PORT=$2; shift 2
  $T 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/tmp.log 2> gpurun_out/tmp.err
  tail -2 gpurun_out/tmp.err; grep '^{' gpurun_out/tmp.log > gpurun_out/$OUT; cut -c1-200 gpurun_out/$OUT; }
OUT=r2_scale_n8.json run 8 29541 --steps 20 --warmup 5 --sustain-s 2
OUT=r2_scale_n4.json run 4 29542 --steps 20 --warmup 5 --sustain-s 2
OUT=r2_cfg2_n8.json run 8 29543 --workload cfg2 --steps 16 --warmup 3 --sustain-s 2
OUT=r2_cfg2_n4.json run 4 29544 --workload cfg2 --steps 16 --warmup 3 --sustain-s 2
$T 120 python -m pytest tests -m gpu -q -x --timeout 100 -k "several_gpus or group" > gpurun_out/r2c13_pytest.log 2>&1; tail -2 gpurun_out/r2c13_pytest.log
python - <<'PY'
import json
for f in ("r2_scale_n8", "r2_scale_n4", "r2_cfg2_n8", "r2_cfg2_n4"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["ms_per_step"], 4), round(d["value"]), round(d["e2e"]["value"]), d["bit_identity"], d["ms_gather_once"], d["ms_per_frame_rank0"], (d["sustained"] or {}).get("ms_per_step"))
    except Exception as e:
        print(f, "unreadable", e)
PY
