python -m pytest tests/test_gpu_primitives.py tests/test_gpu_scheme.py -x -q > gpurun_out/s26_pytest3.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s26_pytest3.log
python bench.py --steps 5 --warmup 3 --no-extras 2>gpurun_out/s26_bench3.err > gpurun_out/s26_bench3.json; python - <<'PY'
import json
d=json.load(open("gpurun_out/s26_bench3.json"))
print("ks/s", d["value"], "parity", d.get("parity_checked"), "resnet20", d["resnet20"]["s_per_image"], d["resnet20"]["numerics"]["max_abs_err"], "r16", d.get("resnet20_logn16",{}).get("s_per_image"), "roofline frac", d["roofline"]["frac"])
PY
for i in 1 2; do python tools/boot_precision.py 2>&1 | tail -1; done
ORION_B200_BC_TC=0 python tools/boot_precision.py 2>&1 | tail -1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:baseconv --csv --log-file gpurun_out/tc_launches_2.csv python tools/boot_precision.py > /dev/null 2>&1
TRACE=resnet20_logn16 python tools/prof_resnet.py 2 --by-op 2>&1 | tail -20 | tee gpurun_out/resnet16_by_op.txt
