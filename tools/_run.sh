mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > gpurun_out/t.log; cat gpurun_out/t.log
python bench.py --steps 50 --warmup 5 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo bench rc=$?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2f_bench_reference_arm.json 2>/dev/null; echo ref rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2f_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_ll.log 2>&1; echo ll rc=$?
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:k_score2ILi0 -s 3 -c 1 -o gpurun_out/prof_r2f python tools/bench_kernel.py --steps 2 > gpurun_out/ncu_r2f.log 2>&1; echo raw rc=$?
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:k_score2ILi1 -s 3 -c 1 -o gpurun_out/prof_r2f_staged python bench.py --steps 2 --warmup 1 --no-cpu --no-criterion > gpurun_out/ncu_r2f_staged.log 2>&1; echo staged rc=$?
python tools/bench_configs.py > gpurun_out/r2f_all_configs.json 2>/dev/null; echo cfg rc=$?
