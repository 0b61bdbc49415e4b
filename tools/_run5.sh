python -m pytest tests/test_gpu_mempressure.py -x -q -s > gpurun_out/mem_pytest.log 2>&1; echo "mem pytest rc=$?"; grep "fatal\|passed\|failed\|peak" gpurun_out/mem_pytest.log | tail -5
tools/profile_ks.sh r2b > gpurun_out/r2b_profile.log 2>&1; echo "profile rc=$?"; tail -3 gpurun_out/r2b_profile.log
