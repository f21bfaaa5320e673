# round-end evidence on one B200: parity tests, smoke, bench (both arms), stage times, ncu launch list + full capture
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/z_tests.log; cat gpurun_out/z_tests.log
python __graft_entry__.py smoke > gpurun_out/z_smoke.log 2>&1; tail -1 gpurun_out/z_smoke.log
python bench.py > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; tail -2 gpurun_out/z_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/z_bench_ref.json 2> gpurun_out/z_bench_ref.err; tail -2 gpurun_out/z_bench_ref.err
python tools/stage_times.py > gpurun_out/z_stage_enc.log 2>&1
python tools/stage_times.py decode > gpurun_out/z_stage_dec.log 2>&1
python tools/stage_times.py c3 > gpurun_out/z_stage_c3.log 2>&1
python tools/bench_configs.py --reps 2 --workers 16 > gpurun_out/z_configs_1gpu.jsonl 2> gpurun_out/z_configs_1gpu.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/z_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/z_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -o gpurun_out/r1_final_full -f python tools/roundtrip_once.py 2160 4096 4 16 smooth 1 > gpurun_out/z_ncu_full.log 2>&1; tail -2 gpurun_out/z_ncu_full.log
python -c "
import json; d=json.load(open('gpurun_out/z_bench.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], 'pipe', d['e2e_pipelined']['value'], 'enc', d['encode_mpix_s'], 'dec', d['decode_mpix_s'], 'roof', d['roofline']['frac'], 'cpu', d['cpu_baseline'])"
cat gpurun_out/z_bench_ref.json | cut -c1-200
