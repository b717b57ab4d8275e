mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/t.log; cat gpurun_out/t.log
python tools/bench_kernel.py; python tools/bench_kernel.py --int64; python tools/bench_kernel.py --mode greedy; python tools/bench_kernel.py --mode scb
python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_f.json'))
print('value',d['value'],'ms',d['ms_per_step']); print(d['staged_refs']); print(d['int64_rows']); print(d['roofline']['issue_frac'], d['roofline']['warp_inst_per_caption'])
PY
