#!/usr/bin/env python
"""Turn an `ncu --set full` capture of one b2col_update (tools/gpujobs/r2z_ncu.sh) into the per-kernel counter summary kept under
profiles/ (run where ncu is installed: python tools/ncu_update_summary.py gpurun_out/r2z_update.ncu-rep > profiles/...txt)."""
import csv
import subprocess
import sys

WANT = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"), ("sm__inst_executed.avg.per_cycle_elapsed", "IPC"),
        ("smsp__inst_executed.sum", "warp instr"), ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads per instr"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "L1 LSU wavefronts %"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %")]


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    total = 0.0
    for r in rows[2:]:
        print(r[idx["Kernel Name"]].split("(")[0])
        for key, label in WANT:
            if key in idx:
                print(f"    {label:28s} {r[idx[key]]:>16s} {units[idx[key]]}")
        stalls = sorted(((float(r[i]), h) for h, i in idx.items() if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") and r[i] not in ("", "n/a")),
                        reverse=True)[:5]
        print("    stalled warps per issue      " + ", ".join(f"{h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')} {v:.2f}" for v, h in stalls))
        total += float(r[idx["gpu__time_duration.sum"]])
    print(f"sum of the captured launches: {total:.1f} us (cold-cache, serialised)")


if __name__ == "__main__":
    main()
