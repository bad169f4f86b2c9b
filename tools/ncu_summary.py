"""Raw-page CSV of an `ncu --set full` report  ->  the small JSON `bench.py` reads for `roofline.traffic`.

    ncu -i prof.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_summary.py raw.csv profiles/rXy_ncu_summary.json

Run it in the tree the capture was made from: the summary is stamped with the hash of the kernel sources
(`kernel_sources_sha`), and `bench.py` ignores summaries whose stamp differs from the sources it is timing.
One entry per kernel (the first captured launch of each name): DRAM bytes read/written, duration, registers,
ALU/FMA pipe and issue-slot activity, executed warp instructions, active warps per cycle.
"""
import csv
import json
import re
import sys


def pick(row, *names):
    for n in names:
        for k, v in row.items():
            if k.endswith(n) and v not in ("", "n/a"):
                return v
    return None


def to_bytes(value, unit):
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(value) * scale.get(unit, 1.0)


def main(src, dst):
    rows = list(csv.reader(open(src)))
    header = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    names, units = rows[header], rows[header + 1]
    unit_of = dict(zip(names, units))
    out = {}
    for r in rows[header + 2:]:
        if len(r) != len(names):
            continue
        row = dict(zip(names, r))
        kernel = re.sub(r"^(void )?(nm03::)?", "", row["Kernel Name"])
        kernel = re.sub(r"\(int\)", "", kernel).split("(")[0]
        if kernel in out:
            continue
        rd_key = next(k for k in names if k.endswith("dram__bytes_read.sum"))
        wr_key = next(k for k in names if k.endswith("dram__bytes_write.sum"))
        dur_key = next(k for k in names if k.endswith("gpu__time_duration.sum"))
        dur = float(row[dur_key]) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit_of[dur_key], 1.0)
        out[kernel] = {
            "dram_read_bytes": to_bytes(row[rd_key], unit_of[rd_key]),
            "dram_write_bytes": to_bytes(row[wr_key], unit_of[wr_key]),
            "duration_us": dur,
            "regs": pick(row, "launch__registers_per_thread"),
            "alu_pipe_pct_active": pick(row, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
            "fma_pipe_pct_active": pick(row, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
                                        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
            "issue_active_pct": pick(row, "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct"),
            "warp_inst": pick(row, "smsp__inst_executed.sum", "sm__inst_executed.sum"),
            "warps_active_per_cycle": pick(row, "sm__warps_active.avg.per_cycle_active"),
        }
    # bench.py quotes a summary only when it was captured from the kernel sources it is timing
    import os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    out["kernel_sources_sha"] = bench.kernel_sources_sha()
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
