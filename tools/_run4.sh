tools/ks_ab.sh ORION_B200_IP_ZFAST 0 1 0 1 2>&1 | tee gpurun_out/zfast_ab.txt
python -m pytest tests/test_gpu_mempressure.py -x -q -s > gpurun_out/mem_pytest.log 2>&1; echo "mem pytest rc=$?"; grep "^\[mem\]\|fatal\|passed\|failed\|peak" gpurun_out/mem_pytest.log | tail -30
python -m pytest tests/test_gpu_primitives.py -x -q 2>&1 | tail -2
