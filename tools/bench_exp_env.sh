# developer tool: bench experimental builds (compiled with -DCST_EXPERIMENT) under CST_EXP_* settings;
# each argument is "lib:VAR=val,VAR=val" (lib = file in cst_captioning_b200/, settings optional)
for cfg in "$@"; do
  lib=${cfg%%:*}; envs=$(echo "${cfg#*:}" | tr ',' ' '); [ "$lib" = "$cfg" ] && envs=""
  env $envs CST_REWARD_LIB=$PWD/cst_captioning_b200/$lib timeout 300 python bench.py --no-cpu --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$cfg', d['ms_per_step'], d['check']['m_score'])"
done
