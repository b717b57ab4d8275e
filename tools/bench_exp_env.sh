# developer tool: bench one experimental build (libcst_exp.so, compiled with -DCST_EXPERIMENT) under
# several CST_EXP_* environment settings; each argument is "VAR=val,VAR=val" ("-" = none)
for cfg in "$@"; do
  envs=$(echo "$cfg" | tr ',' ' '); [ "$cfg" = "-" ] && envs=""
  env $envs CST_REWARD_LIB=$PWD/cst_captioning_b200/libcst_exp.so timeout 300 python bench.py --no-cpu --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$cfg', d['ms_per_step'], d['check']['m_score'])"
done
