#!/bin/bash
# tools/bc_tc_ab.sh -- on the GPU box: the INT8 tensor-core trial of the base-conversion contraction.
# Parity first (bootstrap words vs the oracle with the trial on), then bootstrap latency off/on, then ncu of both kernels.
mkdir -p gpurun_out
ORION_B200_BC_TC=1 python -m pytest tests/test_gpu_scheme.py -q -x -k "bootstrap" > gpurun_out/tc_pytest.log 2>&1; echo "pytest(TC=1) rc=$?"; tail -2 gpurun_out/tc_pytest.log
for rt in 0 1 2 4; do
  if [ $rt = 0 ]; then echo "TC=0"; ORION_B200_BC_TC=0 python tools/boot_precision.py 2>&1 | tail -1
  else echo "TC=1 RT=$rt"; ORION_B200_BC_TC=1 ORION_B200_BC_TC_RT=$rt python tools/boot_precision.py 2>&1 | tail -1; fi
done | tee gpurun_out/tc_boot_ab.txt
for tc in 0 1; do
  ORION_B200_BC_TC=$tc ncu --metrics gpu__time_duration.sum --clock-control none -k regex:baseconv --csv --log-file gpurun_out/tc_launches_$tc.csv \
      python tools/boot_precision.py > gpurun_out/tc_ncu_$tc.log 2>&1
done
ORION_B200_BC_TC=1 ncu --set full --clock-control none --import-source on -k regex:baseconv_tc -s 40 -c 4 -o gpurun_out/tc_bc_tc -f python tools/boot_precision.py > gpurun_out/tc_ncu_full1.log 2>&1
ORION_B200_BC_TC=0 ncu --set full --clock-control none --import-source on -k regex:baseconv_batch -s 60 -c 8 -o gpurun_out/tc_bc_int -f python tools/boot_precision.py > gpurun_out/tc_ncu_full0.log 2>&1
ls -la gpurun_out/
