"""Summarise an .ncu-rep (read with `ncu -i ... --page raw --csv`) into a small text file for profiles/.
python tools/ncu_summary.py <report.ncu-rep> <out.txt>"""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__cycles_elapsed.avg.per_second", "dram__cycles_elapsed.avg.per_second"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    lines = [f"# summary of {rep.split('/')[-1]} (ncu --set full --clock-control none; per-launch values, cold-cache/serialised timing)"]
    for r in rows[2:]:
        lines.append("kernel: " + r[hdr.index("Kernel Name")])
        for w in WANT:
            if w in hdr:
                lines.append("  %-72s %s %s" % (w, r[hdr.index(w)], units[hdr.index(w)]))
        rd, wr = float(r[hdr.index("dram__bytes_read.sum")]), float(r[hdr.index("dram__bytes_write.sum")])
        u = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}
        tot = rd * u[units[hdr.index("dram__bytes_read.sum")]] + wr * u[units[hdr.index("dram__bytes_write.sum")]]
        dur = float(r[hdr.index("gpu__time_duration.sum")]) * {"us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1}[units[hdr.index("gpu__time_duration.sum")]]
        lines.append("  dram_bytes_per_launch (read+write)                                       %.0f  => %.1f GB/s over the launch" % (tot, tot / dur / 1e9))
        st = [(hdr[i], float(r[i])) for i in range(len(hdr)) if "pcsamp_warps_issue_stalled" in hdr[i] and "not_issued" not in hdr[i]]
        tt = sum(v for _, v in st) or 1.0
        lines.append("  warp stall samples: " + ", ".join("%s %.0f%%" % (k.split("stalled_")[1], 100 * v / tt) for k, v in sorted(st, key=lambda kv: -kv[1])[:8]))
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
