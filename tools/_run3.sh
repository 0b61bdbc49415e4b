ORION_B200_BC_TC=1 python -m pytest tests/test_gpu_scheme.py tests/test_gpu_primitives.py tests/test_gpu_models.py -q -x > gpurun_out/tc_pytest2.log 2>&1; echo "pytest(TC=1) rc=$?"; tail -2 gpurun_out/tc_pytest2.log
for z in 3 6 10; do echo "TC=1 RT=2 Z=$z"; ORION_B200_BC_TC=1 ORION_B200_BC_TC_Z=$z python tools/boot_precision.py 2>&1 | tail -1; done | tee gpurun_out/tc_boot_ab2.txt
echo "TC=0"; ORION_B200_BC_TC=0 python tools/boot_precision.py 2>&1 | tail -1 | tee -a gpurun_out/tc_boot_ab2.txt
python -m pytest tests/test_gpu_mempressure.py -x -q -s > gpurun_out/mem_pytest.log 2>&1; echo "mem pytest rc=$?"; tail -5 gpurun_out/mem_pytest.log
for tc in 0 1; do ORION_B200_BC_TC=$tc python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/tc_bench_$tc.json 2>/dev/null; python - <<PY
import json
d=json.load(open("gpurun_out/tc_bench_$tc.json"))
print("TC=$tc resnet20 s/image", d["resnet20"]["s_per_image"], d["resnet20"]["numerics"], "lola", d["lola2"]["s_per_image"], "ks/s", d["value"])
PY
done
ORION_B200_BC_TC=1 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:baseconv --csv --log-file gpurun_out/tc_launches_1c.csv python tools/boot_precision.py > gpurun_out/tc_ncu_1c.log 2>&1
