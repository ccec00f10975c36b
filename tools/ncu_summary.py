"""Condense an `ncu --set full` report into the per-launch table kept under profiles/:
    python tools/ncu_summary.py gpurun_out/r01f_full.ncu-rep profiles/r01f_ncu_full_summary.csv [profiles/traffic.json]
Also prints DRAM bytes per tick (sum over one launch of each tick kernel) for profiles/traffic.json."""
import csv, io, subprocess, sys

KEEP = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]


def main(rep, out, traffic=None):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    cols = [head.index(k) for k in KEEP if k in head]
    cols += [i for i, h in enumerate(head) if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
    with open(out, "w", newline="") as f:
        wr = csv.writer(f)
        wr.writerow([head[i] for i in cols]); wr.writerow([units[i] for i in cols])
        for r in body:
            wr.writerow([r[i] for i in cols])
    name, rd, wrt, dur = head.index("Kernel Name"), head.index("dram__bytes_read.sum"), head.index("dram__bytes_write.sum"), head.index("gpu__time_duration.sum")

    def to_bytes(v, u):
        v = float(v.replace(",", ""))
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    total, per = 0.0, {}
    for r in body:
        b = to_bytes(r[rd], units[rd]) + to_bytes(r[wrt], units[wrt])
        total += b
        per[r[name].split("(")[0]] = {"dram_read_MB": to_bytes(r[rd], units[rd]) / 1e6, "dram_write_MB": to_bytes(r[wrt], units[wrt]) / 1e6,
                                      "duration_us": float(r[dur].replace(",", ""))}
        print(f"{r[name][:40]:40s} {r[dur]:>12s} {units[dur]:3s}  dram {b / 1e6:8.1f} MB")
    print("sum of DRAM bytes over the listed launches:", int(total))
    if traffic:
        import json
        json.dump({"source": f"ncu --set full, {out} (one launch of each tick kernel, 1 M scene, tick 5)",
                   "unit_note": "MB = 1e6 bytes as printed by ncu (Mbyte)", "per_kernel": per, "tick_dram_bytes": total},
                  open(traffic, "w"), indent=1)


if __name__ == "__main__":
    main(*sys.argv[1:4])
