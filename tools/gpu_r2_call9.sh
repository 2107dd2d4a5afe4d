# round 2, call 9 (2 GPUs): group / shared-image tests on two real devices, new tests, bench at N = 2
# (headline weak scaling with the image assembled every frame; configs[2] strong scaling with the
# bit-identity check; the NCCL-gather form beside it)
T="timeout -k 10"
nvidia-smi -L
$T 600 python -m pytest tests -m gpu -q -x --timeout 400 -k "group or full_image or shipped_cube or shared_memory or no_alloc or interleaved" > gpurun_out/r2c9_pytest.log 2>&1; tail -3 gpurun_out/r2c9_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$T 400 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2c9_bench_n2.log 2> gpurun_out/r2c9_bench_n2.err; tail -3 gpurun_out/r2c9_bench_n2.err; grep '^{' gpurun_out/r2c9_bench_n2.log > gpurun_out/r2_scale_n2.json; cut -c1-600 gpurun_out/r2_scale_n2.json
$T 400 $TR bench.py --gpus 2 --workload cfg2 --steps 16 --warmup 3 > gpurun_out/r2c9_cfg2_n2.log 2> gpurun_out/r2c9_cfg2_n2.err; tail -3 gpurun_out/r2c9_cfg2_n2.err; grep '^{' gpurun_out/r2c9_cfg2_n2.log > gpurun_out/r2_cfg2_n2.json; cut -c1-400 gpurun_out/r2_cfg2_n2.json
$T 400 $TR bench.py --gpus 2 --workload cfg2 --steps 16 --warmup 3 --gather nccl --sustain-s 0 > gpurun_out/r2c9_cfg2_n2_nccl.log 2> gpurun_out/r2c9_cfg2_n2_nccl.err; tail -3 gpurun_out/r2c9_cfg2_n2_nccl.err; grep '^{' gpurun_out/r2c9_cfg2_n2_nccl.log > gpurun_out/r2_cfg2_n2_nccl.json; cut -c1-400 gpurun_out/r2_cfg2_n2_nccl.json
$T 300 python bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2c9_ref_n2.json 2>&1; cut -c1-300 gpurun_out/r2c9_ref_n2.json
$T 400 $TR tools/config5.py --mode tiles --frames 4 --out gpurun_out/r2_config5_8k_n2_tiles.json > gpurun_out/r2c9_cfg5.log 2>&1; tail -2 gpurun_out/r2c9_cfg5.log | cut -c1-600
