# decoder variant check: parity + stage times + bench
mkdir -p gpurun_out
python -m pytest tests/test_gpu_decode.py tests/test_gpu_sequence.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/l_tests.log; cat gpurun_out/l_tests.log
python tools/stage_times.py decode 2>&1 | grep -E "t1=" > gpurun_out/l_stage_dec.log; cat gpurun_out/l_stage_dec.log
python bench.py --no-cpu-baseline --steps 4 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(' value %.1f e2e %.1f enc %.1f dec %.1f' % (d['value'], d['e2e']['value'], d['encode_mpix_s'], d['decode_mpix_s']))" > gpurun_out/l_bench.log; cat gpurun_out/l_bench.log
