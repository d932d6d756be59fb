#!/bin/bash
# Round-2 evidence on ONE B200: full GPU suite, bench lines of every config, diagnostics, ncu captures.
# Everything lands in gpurun_out/; the summaries are copied into profiles/ afterwards.
O=gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -8 > $O/r02_pytest_gpu_full.txt
python bench.py > $O/r02_bench_c3_default.json 2> $O/r02_bench_c3_default.err
for w in c1 c2 c4 c5; do python bench.py --workload $w --steps 100 --warmup 10 > $O/r02_bench_$w.json 2>> $O/r02_bench_other.err; done
python bench.py --workload c5 --share-prefix mma --steps 100 --warmup 10 --no-cpu-baseline > $O/r02_bench_c5_group_mma.json 2>> $O/r02_bench_other.err
python bench.py --workload c5 --share-prefix off --steps 100 --warmup 10 --no-cpu-baseline > $O/r02_bench_c5_unshared.json 2>> $O/r02_bench_other.err
python bench.py --workload c2e --steps 100 --warmup 10 > $O/r02_bench_c2e_graph.json 2>> $O/r02_bench_other.err
for n in 2 4 8; do python bench.py --workload c3 --shard-of $n --steps 100 --warmup 10 --no-cpu-baseline > $O/r02_bench_c3_shard${n}_1gpu.json 2>> $O/r02_bench_other.err; done
python bench.py --workload c5 --shard-of 8 --steps 100 --warmup 10 --no-cpu-baseline > $O/r02_bench_c5_shard8_1gpu.json 2>> $O/r02_bench_other.err
python bench.py --impl reference --steps 5 --warmup 1 > $O/r02_bench_c3_reference_arm.json 2>> $O/r02_bench_other.err
: > $O/r02_diag_e2e.jsonl
for w in c3 c4 c2; do python scripts/diag_e2e.py $w >> $O/r02_diag_e2e.jsonl 2>> $O/r02_bench_other.err; done
python scripts/diag_stream_times.py c3 8 > $O/r02_stream_timeline_c3_shard8.json 2>> $O/r02_bench_other.err
python scripts/diag_group_tc5.py > $O/r02_group_tc5_timeline_c5.txt 2>> $O/r02_bench_other.err
# ncu: the exact default command, short
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-neighbours --no-e2e --sustain-s 0 --parity-seqs 0"
$CMD > $O/plain_c3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 3 -c 12 --csv --log-file $O/r02_ncu_launches_c3.csv $CMD > $O/ncu_c3_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paged_decode_split_mma -s 3 -c 1 -f -o $O/r02_split_c3 $CMD > $O/ncu_c3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paged_decode_split_mma -s 3 -c 1 -f -o $O/r02_split_c4 $CMD --workload c4 > $O/ncu_c4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paged_decode_split_mma -s 3 -c 1 -f -o $O/r02_split_c3_shard8 $CMD --shard-of 8 > $O/ncu_c3s8.log 2>&1
tail -3 $O/r02_pytest_gpu_full.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_bench_c*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], "ms", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3) if d.get("roofline") else None,
              "e2e_ms", round(d["e2e"]["ms_per_step"], 4) if d.get("e2e") and "ms_per_step" in d["e2e"] else None,
              "parity", d["parity"]["max_abs_err"] if d.get("parity") else None)
    except Exception as e:
        print(f, "unreadable", e)
PY
