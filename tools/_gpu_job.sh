timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_slab.py -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
python tools/blur_times.py > gpurun_out/blur_times.txt 2>&1; head -3 gpurun_out/blur_times.txt
ncu --set full --import-source on --clock-control none -k regex:blur_fused_unit -c 6 -o gpurun_out/blur_v26 python tools/run_once.py C2_mri_256 1 > gpurun_out/ncu_blur.log 2>&1; tail -2 gpurun_out/ncu_blur.log
