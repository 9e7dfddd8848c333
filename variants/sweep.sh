B="python bench.py --no-cpu --no-e2e"
P='import json,sys; d=json.loads(sys.stdin.readlines()[-1]); print(round(d["value"]/1e6,2), round(d["ms_per_step"],4), d["kernel_ms_per_step"]["k_replan"])'
python -m pytest tests -m gpu -x -q 2>&1 | tail -1
echo -n "dg1: "; $B 2>/dev/null | python -c "$P"
for v in dg2 dg4; do cp variants/$v.so human_robot_gym_b200/libhrg_shield.so; echo -n "$v: "; $B 2>/dev/null | python -c "$P"; done
