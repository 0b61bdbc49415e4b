python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/s26_bench_n2.json 2> gpurun_out/s26_bench_n2.err; echo "bench n2 rc=$?"; tail -c 300 gpurun_out/s26_bench_n2.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/s26_bench_n2.json"))
print("ks/s", d["value"], "parity", d.get("parity_checked"), "resnet20", d["resnet20"]["s_per_image"], "strong", d["resnet20"].get("strong_scaling"), "r16", d.get("resnet20_logn16"))
PY
