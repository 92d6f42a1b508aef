#!/usr/bin/env python
"""Compact per-launch table out of an `ncu --set full ... --page raw --csv` dump of one forward step.

    python tools/ncu_table.py gpurun_out/ncu_step_raw.csv profiles/rNN_ncu_step.csv [--traffic profiles/ncu_traffic.json --batch 256 --model resnet18]

One row per launch: kernel, grid, block, registers, duration, tensor-pipe activity, DRAM bytes read / written, L2 hit
rate, shared-memory wavefronts (LSU) as a share of peak, issue-slot use.  With --traffic it also writes the average
DRAM bytes per conv / FC launch (the `roofline.traffic` figure bench.py reports when the launch count matches).
"""
import argparse
import csv
import json
import re
import subprocess

COLS = [
    ("Kernel Name", "kernel"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"),
    ("gpu__time_duration.sum", "time_us"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct"),
    ("dram__bytes_read.sum", "dram_read"),
    ("dram__bytes_write.sum", "dram_write"),
    ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "lsu_smem_wavefronts_pct"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
]
TO_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TO_US = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3, "second": 1e6}


def short(name):
    m = re.search(r"(\w+_kernel|\w+kernel\w*)", name)
    return m.group(1) if m else name[:40]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("raw")
    ap.add_argument("out")
    ap.add_argument("--traffic")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--model", default="resnet18")
    args = ap.parse_args()
    rows = list(csv.reader(open(args.raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in data:
        if len(r) != len(hdr):
            continue
        rec = {}
        for src, dst in COLS:
            if src not in ix:
                rec[dst] = ""
                continue
            v, u = r[ix[src]], units[ix[src]]
            if dst == "kernel":
                rec[dst] = short(v)
            elif dst in ("dram_read", "dram_write"):
                rec[dst + "_MB"] = round(float(v.replace(",", "")) * TO_BYTES.get(u, 1.0) / 1e6, 3)
            elif dst == "time_us":
                rec[dst] = round(float(v.replace(",", "")) * TO_US.get(u, 1.0), 2)
            else:
                try:
                    rec[dst] = round(float(v.replace(",", "")), 2)
                except ValueError:
                    rec[dst] = v
        out.append(rec)
    keys = list(out[0].keys())
    with open(args.out, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=["launch"] + keys)
        w.writeheader()
        for i, rec in enumerate(out):
            w.writerow(dict(rec, launch=i))
    conv = [r for r in out if re.search(r"stem|conv|umma", r["kernel"])]
    total = sum(r["time_us"] for r in out)
    print("%d launches, %.1f us under ncu (cold cache, serialised); %d conv / FC launches = %.1f %% of that" %
          (len(out), total, len(conv), 100.0 * sum(r["time_us"] for r in conv) / total))
    print("DRAM per step: %.1f MB read + %.1f MB written" % (sum(r["dram_read_MB"] for r in out), sum(r["dram_write_MB"] for r in out)))
    if args.traffic:
        try:
            commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
        except OSError:
            commit = ""
        per = sum(r["dram_read_MB"] + r["dram_write_MB"] for r in conv) * 1e6 / max(1, len(conv))
        json.dump({"model": args.model, "batch": args.batch, "conv_fc_launches": len(conv),
                   "dram_bytes_per_conv_fc_launch": per, "source": "%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, "
                   "average over the %d conv / FC launches of one step; commit %s)" % (args.out, len(conv), commit)},
                  open(args.traffic, "w"), indent=1)
        print("traffic per conv / FC launch: %.1f MB" % (per / 1e6))


if __name__ == "__main__":
    main()
