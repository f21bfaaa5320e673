# 8-GPU (or N-GPU) bench line, as the driver launches it
mkdir -p gpurun_out
N=${1:-8}
nproc; free -g | head -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/m_bench_${N}gpu.json 2> gpurun_out/m_bench_${N}gpu.err
tail -3 gpurun_out/m_bench_${N}gpu.err; python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/m_bench_${N}gpu.json') if l.startswith('{')][-1])
print(d['n_gpus'], 'value %.1f e2e %.1f pipe %.1f enc %.1f dec %.1f' % (d['value'], d['e2e']['value'], d['e2e_pipelined']['value'], d['encode_mpix_s'], d['decode_mpix_s']), d['clocks'])"
