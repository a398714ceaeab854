#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics per captured launch and the
top stall sites of the SASS source page.  Output is the text kept under profiles/.

  python tools/ncu_summary.py gpurun_out/gemv.ncu-rep [--top 25] > profiles/r01_gemv_summary.txt
"""
import csv
import io
import subprocess
import sys

RAW_KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "launch__waves_per_multiprocessor", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_wait_per_warp_active.pct",
]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 25
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print(f"# {rep}: {len(data)} captured launch(es)")
    for r in data:
        name = r[hdr.index("Kernel Name")]
        print(f"\n## {name}")
        for k in RAW_KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:85s} {r[i]:>16s} {units[i]}")
    src = ncu(["-i", rep, "--page", "source", "--csv"])
    blocks = src.split('"Kernel Name",')
    for blk in blocks[1:2]:   # first launch only
        lines = blk.split("\n")
        print(f"\n## stall sites (first launch): {lines[0][:100]}")
        rd = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
        h = rd[0]
        isamp, isrc, iexec = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
        stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
        tot = sum(int(r[isamp] or 0) for r in rd[1:] if len(r) > isamp)
        agg = {}
        for r in rd[1:]:
            if len(r) <= isamp:
                continue
            for i in stall_cols:
                agg[h[i]] = agg.get(h[i], 0) + int(r[i] or 0)
        print(f"  total samples {tot}; by reason: " + ", ".join(f"{k}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
        ranked = sorted((r for r in rd[1:] if len(r) > isamp), key=lambda r: -int(r[isamp] or 0))[:top]
        for r in ranked:
            why = sorted(((int(r[i] or 0), h[i]) for i in stall_cols), reverse=True)[:2]
            print(f"  {int(r[isamp]):6d} ({100.0 * int(r[isamp]) / max(tot, 1):4.1f}%) exec={r[iexec]:>7s}  {r[isrc].strip()[:70]:70s} {why[0][1]}={why[0][0]} {why[1][1]}={why[1][0]}")


if __name__ == "__main__":
    main()
