#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into the few numbers DESIGN.md quotes.

    python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep [units_per_launch] > profiles/<name>.md

Prints, per profiled launch: duration, DRAM bytes read/written (the "traffic" of bench.py's
roofline), DRAM and issue utilisation, registers, occupancy limits, shared-memory bank conflicts,
and — from the SASS source page of the first launch — the opcode mix and the top stall sites.
"""
import collections
import csv
import io
import subprocess
import sys

RAW = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum", "lts__t_sector_hit_rate.pct",
    "sm__cycles_elapsed.avg.per_second",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
    "smsp__sass_inst_executed_op_tmem_ldt.sum", "smsp__sass_inst_executed_op_tmem_stt.sum",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
]


def ncu(rep, *args):
    out = subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 else None
    rows = ncu(rep, "--page", "raw", "--csv")
    hdr, unit = rows[0], rows[1]
    print(f"# ncu summary of `{rep.split('/')[-1]}`\n")
    for r in rows[2:]:
        print(f"## launch {r[hdr.index('ID')]}: `{r[hdr.index('Kernel Name')][:100]}`  grid "
              f"{r[hdr.index('Grid Size')]} block {r[hdr.index('Block Size')]}\n")
        print("| metric | value | unit |\n|---|---|---|")
        for m in RAW:
            if m in hdr:
                print(f"| {m} | {r[hdr.index(m)]} | {unit[hdr.index(m)]} |")
        print()
    src = ncu(rep, "--page", "source", "--csv", "--print-source", "sass")
    if len(src) < 3:
        return
    h = src[1]
    data = src[2:]
    for k, r in enumerate(data):
        if r and r[0] == "Kernel Name":
            data = data[:k]
            break
    iS, iE, iN = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    ops, total = collections.Counter(), 0
    for r in data:
        toks = r[iS].split()
        op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
        ops[op] += int(r[iE])
        total += int(r[iE])
    print("## SASS opcode mix (warp instructions executed, first launch)\n")
    print("| opcode | executed | share |" + (" per unit |" if units else "") + "\n|---|---|---|" + ("---|" if units else ""))
    for op, c in ops.most_common(18):
        print(f"| {op} | {c} | {100 * c / total:.1f}% |" + (f" {c / units:.1f} |" if units else ""))
    print(f"| total | {total} | 100% |" + (f" {total / units:.1f} |" if units else ""))
    stalls = collections.Counter()
    cols = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
    for r in data:
        for i in cols:
            stalls[h[i]] += int(r[i] or 0)
    tot = sum(stalls.values()) or 1
    print("\n## warp stall samples (all)\n\n| reason | samples | share |\n|---|---|---|")
    for n, c in stalls.most_common(8):
        print(f"| {n} | {c} | {100 * c / tot:.1f}% |")
    print("\n## hottest instructions by samples\n\n| samples | executed | SASS |\n|---|---|---|")
    for r in sorted(data, key=lambda r: -int(r[iN] or 0))[:12]:
        print(f"| {r[iN]} | {r[iE]} | `{r[iS].strip()[:90]}` |")


if __name__ == "__main__":
    main()
