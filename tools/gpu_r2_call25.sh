# round 2, call 25: the GPU suite at HEAD (without the 26 s 10 M-triangle case: the tree is hash-identical)
timeout -k 5 45 python -m pytest tests -m gpu -q -x --timeout 40 -k "not config4" 2>&1 | tail -4 > gpurun_out/r2_call25_pytest.log
cat gpurun_out/r2_call25_pytest.log
