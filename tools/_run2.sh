ORION_B200_BC_TC=1 python -m pytest tests/test_gpu_scheme.py -q -x -k "bootstrap" > gpurun_out/tc_pytest.log 2>&1; echo "pytest(TC=1) rc=$?"; tail -2 gpurun_out/tc_pytest.log
for rt in 0 1 2 4; do
  if [ $rt = 0 ]; then echo "TC=0"; ORION_B200_BC_TC=0 python tools/boot_precision.py 2>&1 | tail -1
  else echo "TC=1 RT=$rt"; ORION_B200_BC_TC=1 ORION_B200_BC_TC_RT=$rt python tools/boot_precision.py 2>&1 | tail -1; fi
done | tee gpurun_out/tc_boot_ab.txt
python -m pytest tests/test_gpu_mempressure.py -x -q -s > gpurun_out/mem_pytest.log 2>&1; echo "mem pytest rc=$?"; tail -5 gpurun_out/mem_pytest.log
ORION_B200_DEBUG=1 timeout 900 python tools/resnet16.py 2 > gpurun_out/resnet16.json 2> gpurun_out/resnet16.err; echo "resnet16 rc=$?"; grep -v "^\[boot\]" gpurun_out/resnet16.err | tail -25; cat gpurun_out/resnet16.json
