python -m pytest tests -m gpu -x -q > gpurun_out/s26_pytest2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s26_pytest2.log
python bench.py > gpurun_out/s26_bench2.json 2> gpurun_out/s26_bench2.err; echo "bench rc=$?"; tail -c 400 gpurun_out/s26_bench2.err
python tools/batch_images.py resnet20 1 2 4 8 2>&1 | tee gpurun_out/batch_images.txt
