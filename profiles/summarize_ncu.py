#!/usr/bin/env python
"""Turn ncu outputs brought back in gpurun_out/ into the small text/CSV/JSON summaries kept
under profiles/ (the .ncu-rep files themselves are scratch and not committed).

    python profiles/summarize_ncu.py <launches.csv> <report.ncu-rep> <tag>

Writes profiles/<tag>_launches.csv (per-kernel average duration and share of the step),
profiles/<tag>_ncu_full.csv (selected `--set full` metrics per kernel) and updates
profiles/traffic.json (dram bytes read+written per launch, keyed by bench.py kernel name).
"""
import csv
import json
import subprocess
import sys
from collections import defaultdict
from pathlib import Path

HERE = Path(__file__).resolve().parent
BENCH_NAME = {"letterbox_kernel": "K1_letterbox_normalize_chw", "decode_filter": "K2K3_decode_filter_nms",
              "crop_resize_kernel": "K4_roi_crop_resize_normalize",
              "resize_tiles_kernel": "K4_roi_crop_resize_normalize"}
KEEP = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "launch__waves_per_multiprocessor", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]


def main(launches: str, report: str, tag: str) -> None:
    rows = [r for r in csv.reader(l for l in open(launches) if l.startswith('"'))]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    per = defaultdict(list)
    for r in rows[1:]:
        per[r[ki].split("(")[0].replace("void ", "").replace("ia::", "")].append(float(r[vi].replace(",", "")))
    total = sum(sum(v) / len(v) for v in per.values())
    with open(HERE / f"{tag}_launches.csv", "w") as fh:
        fh.write("kernel,launches,avg_us,share_of_step\n")
        for k, v in per.items():
            fh.write(f"{k},{len(v)},{sum(v) / len(v) / 1e3:.2f},{sum(v) / len(v) / total:.4f}\n")

    raw = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    keep = [k for k in KEEP if k in hdr]
    traffic_path = HERE / "traffic.json"
    traffic = json.loads(traffic_path.read_text()) if traffic_path.exists() else {}
    with open(HERE / f"{tag}_ncu_full.csv", "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(keep)
        w.writerow([units[hdr.index(k)] for k in keep])
        for r in rows[2:]:
            w.writerow([r[hdr.index(k)] for k in keep])
            name = r[hdr.index("Kernel Name")]
            for frag, bench_name in BENCH_NAME.items():
                if frag in name:
                    rd, wr = (float(r[hdr.index(k)]) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
                    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}
                    rd *= scale[units[hdr.index("dram__bytes_read.sum")]]
                    wr *= scale[units[hdr.index("dram__bytes_write.sum")]]
                    traffic[bench_name] = int(rd + wr)
    traffic["_source"] = f"{tag}: ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch (B=256, C5)"
    traffic_path.write_text(json.dumps(traffic, indent=1) + "\n")


if __name__ == "__main__":
    main(*sys.argv[1:4])
