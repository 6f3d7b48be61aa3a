python -m pytest tests/test_gpu_scaff.py tests/test_gpu_sharding.py -m gpu -x -q 2>&1 | tail -3
for tmv in -1 2; do
echo "== LXY_TABLE_MODE=$tmv"
for w in cfg4m cfg3 cfg2; do
  LXY_TABLE_MODE=$tmv python bench.py --workload $w --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$w',d['value'],d['ms_per_step'],d['kernel_ms'])"
done; done
