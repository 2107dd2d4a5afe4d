# round 2, call 12 (2 GPUs): DMA-copy assembly at N = 2 (bit identity through cfg2's check), raytrace() overload over a group
T="timeout -k 10"
$T 300 python -m pytest tests -m gpu -q -x --timeout 200 -k "several_gpus or source_compatible or strips_scattered or group" > gpurun_out/r2c12_pytest.log 2>&1; tail -3 gpurun_out/r2c12_pytest.log
run() { N=$1; PORT=$2; shift 2
  $T 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/tmp.log 2> gpurun_out/tmp.err
  tail -2 gpurun_out/tmp.err; grep '^{' gpurun_out/tmp.log > gpurun_out/$OUT; cut -c1-200 gpurun_out/$OUT; }
OUT=r2_scale_n2.json run 2 29531 --steps 20 --warmup 5
OUT=r2_cfg2_n2.json run 2 29532 --workload cfg2 --steps 16 --warmup 3
OUT=r2_cfg2_n2_store.json run 2 29533 --workload cfg2 --steps 16 --warmup 3 --gather store --sustain-s 0
python - <<'PY'
import json, glob
for f in ("r2_scale_n2", "r2_cfg2_n2", "r2_cfg2_n2_store"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["ms_per_step"], 4), round(d["value"]), round(d["e2e"]["value"]), d["bit_identity"], d["ms_gather_once"], d["ms_per_frame_rank0"], (d["sustained"] or {}).get("ms_per_step"))
    except Exception as e:
        print(f, "unreadable", e)
PY
