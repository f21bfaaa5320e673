# decoder check after a Tier-1 decoder change + C3 stage times
mkdir -p gpurun_out
python -m pytest tests/test_gpu_decode.py tests/test_gpu_configs.py tests/test_gpu_sequence.py -m gpu -x -q 2>&1 | tail -4 > gpurun_out/g_tests.log; cat gpurun_out/g_tests.log
python tools/stage_times.py decode > gpurun_out/g_stage_dec.log 2>&1; cat gpurun_out/g_stage_dec.log
python tools/stage_times.py c3 > gpurun_out/g_stage_c3.log 2>&1; cat gpurun_out/g_stage_c3.log
