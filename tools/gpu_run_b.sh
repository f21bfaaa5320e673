# encoder checks after a Tier-1 encoder change: parity tests, then the MQ lanes-per-warp / long-block sweep
mkdir -p gpurun_out
python -m pytest tests/test_gpu_stages.py tests/test_gpu_encode.py tests/test_gpu_configs.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/b_tests.log
cat gpurun_out/b_tests.log
for nl in 0 148 592 1184; do for l in 2 4 8; do echo "NLONG $nl LANES $l"; J2KCUDA_MQ_NLONG=$nl J2KCUDA_MQ_LANES=$l python tools/stage_times.py dwt 2>&1 | grep -E "t1mq"; done; done > gpurun_out/b_lanes.log 2>&1
cat gpurun_out/b_lanes.log
python tools/stage_times.py 2>&1 | head -12 > gpurun_out/b_stage.log; cat gpurun_out/b_stage.log
