"""Turn the ncu --set full capture of k_trace_lbvh into profiles/r2_traffic.json (read by bench.py as roofline.traffic).

    on the GPU box (tools/gpujobs/r2f_ncu.sh):
        ncu --set full --clock-control none --import-source on -k regex:k_trace_lbvh -s 3 -c 1 -o gpurun_out/r2_trace \
            python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras
    here:
        python tools/capture_traffic.py gpurun_out/r2_trace.ncu-rep [--colliders 1000000 --rays 16777216]

Records dram__bytes_read.sum + dram__bytes_write.sum of that ONE launch together with the hash of the sources the
kernel is compiled from (bench.source_hash): bench.py drops the figure when the sources have changed since.
"""
from __future__ import annotations

import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def raw_page(rep: str) -> list[dict]:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    head = rows[0]
    return [dict(zip(head, r)) for r in rows[2:] if len(r) == len(head)]  # rows[1] = units


def num(x: str) -> float:
    return float(x.replace(",", ""))


def units_of(rep: str) -> dict:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(txt)))
    return dict(zip(rr[0], rr[1]))


def config5(args, rows, bench) -> None:
    """dram bytes of the 10M-collider trace (the 640 MB tree does not fit L2: this one IS HBM traffic) -> profiles/r2_traffic_config5.json"""
    if len(rows) < 7:
        raise SystemExit(f"expected 7 captured launches (aimed: 0, random: 6), found {len(rows)}")
    units = units_of(args.report)
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tscale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
    out = {"kernel": "k_trace_lbvh", "workload": "config5", "colliders": 10_000_000, "rays": 16 * 1024 * 1024, "source_hash": bench.source_hash(),
           "report": os.path.basename(args.report),
           "source": "ncu --set full --clock-control none -k regex:k_trace_lbvh -s 4 -c 7 python tools/bench_extras.py config5 (launch 0: aimed rays, launch 6: random rays)"}
    for name, r in (("aimed", rows[0]), ("random", rows[6])):
        rd = num(r["dram__bytes_read.sum"]) * scale.get(units.get("dram__bytes_read.sum", "byte"), 1.0)
        wr = num(r["dram__bytes_write.sum"]) * scale.get(units.get("dram__bytes_write.sum", "byte"), 1.0)
        out[name] = {"dram_bytes_read": int(rd), "dram_bytes_write": int(wr),
                     "duration_ms": num(r["gpu__time_duration.sum"]) * tscale.get(units.get("gpu__time_duration.sum", "ns"), 1e-6),
                     "lts_sector_hit_rate_pct": num(r.get("lts__t_sector_hit_rate.pct", "nan"))}
    path = args.out or os.path.join(ROOT, "profiles", "r2_traffic_config5.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--workload", default="config3")
    ap.add_argument("--colliders", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=16 * 1024 * 1024)
    ap.add_argument("--out", default=None)
    ap.add_argument("--config5", action="store_true", help="report of `python tools/bench_extras.py config5` captured with -s 4 -c 7: launch 0 = aimed rays, launch 6 = random rays")
    args = ap.parse_args()
    import bench

    rows = [r for r in raw_page(args.report) if "k_trace_lbvh" in r.get("Kernel Name", "")]
    if not rows:
        raise SystemExit("no k_trace_lbvh launch in the report")
    if args.config5:
        return config5(args, rows, bench)
    args.out = args.out or os.path.join(ROOT, "profiles", "r2_traffic.json")
    r = rows[-1]
    units = {}
    rd, wr = num(r["dram__bytes_read.sum"]), num(r["dram__bytes_write.sum"])
    # ncu prints these in the unit of the units row (byte / Kbyte / Mbyte / Gbyte): re-read it
    txt = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(txt)))
    for name, unit in zip(rr[0], rr[1]):
        units[name] = unit
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rd *= scale.get(units.get("dram__bytes_read.sum", "byte"), 1.0)
    wr *= scale.get(units.get("dram__bytes_write.sum", "byte"), 1.0)
    dur = num(r["gpu__time_duration.sum"])
    dur_ms = dur * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(units.get("gpu__time_duration.sum", "ns"), 1e-6)
    out = {"kernel": "k_trace_lbvh", "workload": args.workload, "colliders": args.colliders, "rays": args.rays,
           "dram_bytes_read": int(rd), "dram_bytes_write": int(wr), "duration_ms": dur_ms,
           "lts_sector_hit_rate_pct": num(r.get("lts__t_sector_hit_rate.pct", "nan")),
           "l1tex_lsu_wavefront_pct": num(r.get("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "nan")),
           "registers_per_thread": int(num(r.get("launch__registers_per_thread", "0"))),
           "source_hash": bench.source_hash(), "report": os.path.basename(args.report),
           "source": "ncu --set full --clock-control none --import-source on -k regex:k_trace_lbvh -s 3 -c 1 python bench.py --steps 1 --warmup 3 "
                     "--no-cpu-baseline --no-extras; dram__bytes_read.sum + dram__bytes_write.sum of that one launch"}
    json.dump(out, open(args.out, "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
