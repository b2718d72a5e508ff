#!/usr/bin/env python
"""Reads an .ncu-rep (ncu --set full --import-source on, kernels built with -lineinfo) here on the CPU box and prints
(1) the headline counters and (2) warp-instruction share + average active lanes per device function of
snail_b200/csrc/snl_device.cuh (line ranges are looked up from the source, so it follows edits).

    python tools/ncu_breakdown.py gpurun_out/prof.ncu-rep [--sass PATTERN]
"""
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEAD = ["gpu__time_duration.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def function_ranges(path):
    """[(name, first_line, last_line)] of the top-level __device__ functions / structs of a source file."""
    lines = open(path).read().split("\n")
    out, name, start = [], None, 0
    for n, l in enumerate(lines, 1):
        m = re.match(r"^(?:template.*\n)?__device__ __forceinline__ \S+ \*?(\w+)\(", l) or \
            re.match(r"^__device__ __forceinline__ .*?(\w+)\(", l)
        if m:
            if name:
                out.append((name, start, n - 1))
            name, start = m.group(1), n
    if name:
        out.append((name, start, len(lines)))
    return out


def main():
    rep = sys.argv[1]
    raw = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
    for h, u, v in zip(raw[0], raw[1], raw[2]):
        if h in HEAD:
            print(f"{h:90s} {v} {u}")
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--csv", "--print-source", "cuda,sass"))))
    hdr, cur, fname, seen = None, None, None, {}
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            fname = os.path.basename(r[1])
            continue
        if len(r) > 3 and r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or len(r) < len(hdr) - 2:
            continue
        if r[0] != "":
            cur = (fname, int(r[0]))
            continue
        if r[2] in ("...", "-", ""):
            continue
        d = dict(zip(hdr, r))
        seen[r[2]] = (cur, r[3].strip(), int(d["Instructions Executed"]), int(d["Thread Instructions Executed"]))
    ranges = {f: function_ranges(os.path.join(ROOT, "snail_b200", "csrc", f)) for f in ("snl_device.cuh",)}
    tot = sum(o[2] for o in seen.values())
    agg = {}
    for (f, line), sass, w, t in seen.values():
        name = f
        for n, a, b in ranges.get(f, []):
            if a <= line <= b:
                name = n
        e = agg.setdefault(name, [0, 0])
        e[0] += w
        e[1] += t
    print(f"\nwarp instructions: {tot}")
    for k, (w, t) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"  {k:24s} {w / tot * 100:5.1f}% of warp-inst   avg active lanes {t / max(w, 1):5.1f}")
    # dynamic opcode mix by issue pipe (B300_MICROARCH.md: FFMA/FMUL/FADD/IMAD on the fma pipe, FMNMX/IADD3/LOP3/SEL/ISETP/
    # FSETP/MOV... on the alu pipe, each one warp instruction per 2 cycles per SMSP)
    fma = {"FFMA", "FMUL", "FADD", "IMAD", "HFMA2", "FFMA2", "FMUL2", "FADD2"}
    alu = {"FMNMX", "FMNMX3", "IADD3", "LOP3", "SHF", "PRMT", "ISETP", "FSETP", "SEL", "FSEL", "MOV", "PLOP3", "IABS", "LEA",
           "VIADD", "VIMNMX", "IMNMX", "I2FP", "VOTE", "FCHK", "IADD", "VIMNMX3", "UISETP", "VIADDMNMX"}
    ops, pipes = {}, {}
    for (_, sass, w, _t) in seen.values():
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", sass)
        if not m:
            continue
        op = m.group(2)
        ops[op] = ops.get(op, 0) + w
        pipe = "fma" if op in fma else "alu" if op in alu else "lsu" if re.match(r"LD|ST|ATOM|RED", op) else \
            "control" if op in ("BRA", "BSSY", "BSYNC", "WARPSYNC", "CALL", "RET", "EXIT", "BAR", "BREAK", "NOP") else "other"
        pipes[pipe] = pipes.get(pipe, 0) + w
    print("\nissue pipes: " + "  ".join(f"{k} {v / tot * 100:.1f}%" for k, v in sorted(pipes.items(), key=lambda kv: -kv[1])))
    print("opcodes:     " + "  ".join(f"{k} {v / tot * 100:.1f}%" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:16]))
    if "--sass" in sys.argv:
        pat = sys.argv[sys.argv.index("--sass") + 1]
        for addr, ((f, line), sass, w, t) in sorted(seen.items(), key=lambda kv: int(kv[0], 16)):
            if re.search(pat, f"{f}:{line} {sass}"):
                print(f"{f}:{line:<4d} {addr[-5:]} {sass[:64]:64s} {w:>10d} {t / max(w, 1):5.1f}")


if __name__ == "__main__":
    main()
