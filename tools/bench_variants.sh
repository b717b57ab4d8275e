# developer tool: bench experimental builds (cst_captioning_b200/libcst_var_*.so) through CST_REWARD_LIB
for f in cst_captioning_b200/libcst_var_*.so; do
  CST_REWARD_LIB=$PWD/$f python bench.py --no-cpu --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$f', d['ms_per_step'], d['check']['m_score'])"
done
