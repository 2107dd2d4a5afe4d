# round 2, call 23 (what is left of the budget): the soups through the device walk, the builder-dependent parity tests
T="timeout -k 5"
$T 100 python -m pytest tests -m gpu -q -x --timeout 90 -k "soups or large_mesh or hits_bit_identical" -rxX 2>&1 | tail -15 > gpurun_out/r2_call23_pytest.log
cat gpurun_out/r2_call23_pytest.log
