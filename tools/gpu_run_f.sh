# bench line incl. the pipelined arm (1 GPU)
mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; tail -3 gpurun_out/f_bench.err
python -c "
import json; d=json.load(open('gpurun_out/f_bench.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], 'pipelined', d['e2e_pipelined'], 'enc', d['encode_mpix_s'], 'dec', d['decode_mpix_s'])"
