python tools/boot16.py 16 16384 3 | tee gpurun_out/boot16.txt
S=$(grep -o "launches before [0-9]*" gpurun_out/boot16.txt | grep -o "[0-9]*"); C=$(grep -o "per bootstrap [0-9]*" gpurun_out/boot16.txt | grep -o "[0-9]*")
echo "skip $S count $C"
ncu --metrics gpu__time_duration.sum --clock-control none -s $S -c $C --csv --log-file gpurun_out/boot16_launches.csv python tools/boot16.py 16 16384 1 > gpurun_out/boot16_ncu.log 2>&1
python tools/agg_launches.py gpurun_out/boot16_launches.csv 30
