mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/o_tests.log; cat gpurun_out/o_tests.log
python tools/stage_times.py 2>&1 | head -9 > gpurun_out/o_stage_enc.log; cat gpurun_out/o_stage_enc.log
python tools/stage_times.py decode > gpurun_out/o_stage_dec.log 2>&1; cat gpurun_out/o_stage_dec.log
