"""Turns the ncu artefacts in gpurun_out/ into the tracked summaries under profiles/.
  python tools/ncu_summary.py launches gpurun_out/r01_launches_fp32.csv profiles/r01_launches_fp32.md
  python tools/ncu_summary.py full gpurun_out/r01_bps.ncu-rep profiles/r01_bps_ncu_full.md
  python tools/ncu_summary.py traffic gpurun_out/r02_chain.ncu-rep profiles/ncu_traffic.json   (merged into the JSON bench.py reads)
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_umma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max", "lts__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct"]


def launches(src, dst):
    lines = [l for l in open(src) if not l.startswith("==")]
    agg = collections.OrderedDict()
    tot = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = row["Kernel Name"].split("(")[0][:70] + " grid=" + row["Grid Size"]
        v = float(row["Metric Value"].replace(",", ""))
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    with open(dst, "w") as f:
        f.write(f"# ncu launch list summary of `{src}`\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare SHARES).\n\n")
        f.write("| kernel | launches | total ms | share |\n|---|---:|---:|---:|\n")
        for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write(f"| `{k}` | {n} | {t / 1e6:.3f} | {100 * t / tot:.1f}% |\n")
        f.write(f"| **total** | {sum(v[0] for v in agg.values())} | {tot / 1e6:.3f} | 100% |\n")


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write(f"# ncu --set full summary of `{src}`\n\n")
        for r in rows[2:]:
            name = r[hdr.index("Kernel Name")]
            f.write(f"## `{name[:110]}`\n\n| metric | unit | value |\n|---|---|---:|\n")
            for i, h in enumerate(hdr):
                if h in KEYS:
                    f.write(f"| {h} | {units[i]} | {r[i]} |\n")
            f.write("\n")


def traffic(src, dst):
    """Per-kernel DRAM traffic / duration / pipe activity of one `ncu --set full` capture, merged into the JSON that
    bench.py's `roofline.traffic` reads (so the number in the bench line is reproducible from a tracked file)."""
    import json
    import os
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    db = json.load(open(dst)) if os.path.exists(dst) else {}

    def val(r, key, want_unit=None):
        i = hdr.index(key)
        v = float(r[i].replace(",", ""))
        u = units[i].lower()
        if want_unit == "mb":
            v *= {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}[u]
        if want_unit == "ms":
            v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3, "s": 1e3}[u]
        return v
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "").strip()
        rec = {"dram_read_mb": val(r, "dram__bytes_read.sum", "mb"), "dram_write_mb": val(r, "dram__bytes_write.sum", "mb"),
               "time_ms": val(r, "gpu__time_duration.sum", "ms"), "grid": r[hdr.index("launch__grid_size")], "source": os.path.basename(src)}
        for k, short in (("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_active_pct"),
                         ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_active_pct"),
                         ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
                         ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct")):
            if k in hdr:
                rec[short] = val(r, k)
        rec["dram_total_mb"] = rec["dram_read_mb"] + rec["dram_write_mb"]
        db[name] = rec
    json.dump(db, open(dst, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    {"launches": launches, "full": full, "traffic": traffic}[sys.argv[1]](sys.argv[2], sys.argv[3])
