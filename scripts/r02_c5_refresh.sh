#!/bin/bash
# Refresh of the C5 evidence after the group kernel's role split: suite, bench lines, role / item timeline.
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > $O/r02_pytest_gpu_full.txt
python bench.py --workload c5 --steps 100 --warmup 10 > $O/r02_bench_c5.json 2> $O/r02_bench_c5.err
python bench.py --workload c5 --shard-of 8 --steps 100 --warmup 10 --no-cpu-baseline > $O/r02_bench_c5_shard8_1gpu.json 2>> $O/r02_bench_c5.err
python bench.py --workload c5 --share-prefix mma --steps 100 --warmup 10 --no-cpu-baseline > $O/r02_bench_c5_group_mma.json 2>> $O/r02_bench_c5.err
python scripts/diag_group_tc5.py > $O/r02_group_tc5_timeline_c5.txt 2>> $O/r02_bench_c5.err
tail -2 $O/r02_pytest_gpu_full.txt
python - <<'PY'
import json
for f in ("r02_bench_c5", "r02_bench_c5_shard8_1gpu", "r02_bench_c5_group_mma"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, round(d["ms_per_step"], 4), round(d["roofline"]["frac"], 4), d["roofline"]["traffic"], d["gpu_launches"], d["parity"]["max_abs_err"] if d.get("parity") else None,
          round(d["e2e"]["ms_per_step"], 4) if d.get("e2e") else None)
PY
