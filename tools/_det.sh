cd /root/repo
env REPS=10 timeout 200 python tools/det_stress.py full 2>&1 | tail -1
env REPS=12 B2_CHAIN=0 GRAPHS=0 timeout 200 python tools/det_stress.py tiny 2>&1 | tail -1
env REPS=12 GRAPHS=0 timeout 200 python tools/det_stress.py tiny 2>&1 | tail -1
env REPS=8 B2_CHAIN=0 timeout 200 python tools/det_stress.py full 2>&1 | tail -1
STAGES="tc attn engine" bash tools/gpu_ci.sh 2>&1 | grep -E "^===|rc=|passed|failed|C1 greedy|^E  "
python bench.py --value-only 2>&1 | tail -1
B2_CHAIN=0 python bench.py --value-only 2>&1 | tail -1
