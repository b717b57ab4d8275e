mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_parity.py::test_staged_refs_give_the_same_bits_as_raw_rows 2>&1 | tail -5 > gpurun_out/t2.log
for v in reward var_B var_C var_D; do CST_REWARD_LIB=cst_captioning_b200/libcst_$v.so python tools/bench_kernel.py >> gpurun_out/var1.jsonl 2>>gpurun_out/var1.err; done
CST_REWARD_LIB=cst_captioning_b200/libcst_var_B.so python -m pytest tests/test_gpu_parity.py -q -x -k "cfg or random or maximum or long_rows or ragged or edge" 2>&1 | tail -3 >> gpurun_out/t2.log
python tools/bench_kernel.py --int64 >> gpurun_out/var1.jsonl
python tools/bench_kernel.py --mode greedy >> gpurun_out/var1.jsonl
CST_KERNEL=1 python tools/bench_kernel.py --mode greedy >> gpurun_out/var1.jsonl
cat gpurun_out/t2.log; cat gpurun_out/var1.jsonl
