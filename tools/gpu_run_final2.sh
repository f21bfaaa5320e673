# round-end numbers with the final host code (kernels unchanged since the ncu captures of tools/gpu_run_final.sh)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/y_tests.log; cat gpurun_out/y_tests.log
python __graft_entry__.py smoke > gpurun_out/y_smoke.log 2>&1; tail -1 gpurun_out/y_smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/y_bench_ref.json 2> gpurun_out/y_bench_ref.err
python bench.py > gpurun_out/y_bench.json 2> gpurun_out/y_bench.err; tail -2 gpurun_out/y_bench.err
python tools/stage_times.py > gpurun_out/y_stage_enc.log 2>&1
python tools/stage_times.py decode > gpurun_out/y_stage_dec.log 2>&1
python tools/stage_times.py c3 > gpurun_out/y_stage_c3.log 2>&1
python tools/bench_configs.py --reps 2 --workers 16 > gpurun_out/y_configs_1gpu.jsonl 2> gpurun_out/y_configs_1gpu.err
python -c "
import json; d=json.load(open('gpurun_out/y_bench.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], 'pipe', d['e2e_pipelined']['value'], 'enc', d['encode_mpix_s'], 'dec', d['decode_mpix_s'], 'roof', d['roofline']['frac'], 'cpu', d['cpu_baseline']['value'])"
cut -c1-160 gpurun_out/y_bench_ref.json; cut -c1-200 gpurun_out/y_configs_1gpu.jsonl
