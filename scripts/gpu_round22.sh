mkdir -p gpurun_out
for i in 1 2 3; do timeout 200 python -m pytest tests/test_gpu_overlap.py -x -q > gpurun_out/pytest_overlap_$i.log 2>&1; echo "overlap run $i rc=$?"; tail -2 gpurun_out/pytest_overlap_$i.log | cut -c1-200; done
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-300
