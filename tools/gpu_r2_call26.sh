# round 2, call 26 (last seconds of the budget): the 10 M-triangle parity case on the restructured builder
timeout -k 3 29 python -m pytest tests -m gpu -q -x -k "config4" --durations=3 2>&1 | tail -6 > gpurun_out/r2_call26_pytest.log
cat gpurun_out/r2_call26_pytest.log
