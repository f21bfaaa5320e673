mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/n_tests.log; cat gpurun_out/n_tests.log
python bench.py --no-cpu-baseline --steps 4 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(' value %.1f e2e %.1f enc %.1f dec %.1f pipe %.1f' % (d['value'], d['e2e']['value'], d['encode_mpix_s'], d['decode_mpix_s'], d['e2e_pipelined']['value']))" > gpurun_out/n_bench.log; cat gpurun_out/n_bench.log
