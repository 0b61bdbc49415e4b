#!/bin/bash
# tools/profile_ks.sh TAG -- on the GPU box: the key-switch bench without ncu first (must exit 0), then the ncu launch list of
# the same command and one --set full capture of the dominant kernel.  Outputs under gpurun_out/.
TAG=${1:-r2}
set -e
python bench.py --ks-only --steps 2 --warmup 3 > gpurun_out/${TAG}_ks_plain.json 2> gpurun_out/${TAG}_ks_plain.err
# launch list: skip context set-up launches (evk_to_device, uploads), take one full step's worth (768 ks / 8 per group * 11 launches ~ 1056) -> 1200
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 1200 --csv --log-file gpurun_out/${TAG}_launches_ks.csv \
    python bench.py --ks-only --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ntt2_ip_kernel -s 6 -c 2 -o gpurun_out/${TAG}_ks_ip -f \
    python bench.py --ks-only --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/${TAG}_*
