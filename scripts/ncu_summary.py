"""Condense `ncu --page raw --csv` exports into the few columns the profiles/ tables quote.

    python scripts/ncu_summary.py gpurun_out/ncu_decode_fp16_t1.raw.csv [...] > profiles/<name>.csv
"""
import csv
import sys

COLS = [("kernel", "Kernel Name"), ("grid", "launch__grid_size"), ("block", "launch__block_size"),
        ("regs", "launch__registers_per_thread"), ("dyn_smem_KB", "launch__shared_mem_per_block_dynamic"),
        ("occupancy_limit_blocks", "launch__occupancy_limit_shared_mem"),
        ("duration_us", "gpu__time_duration.sum"), ("dram_read_MB", "dram__bytes_read.sum"),
        ("dram_write_MB", "dram__bytes_write.sum"),
        ("dram_pct_of_peak", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        ("tensor_pipe_pct_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
        ("tensor_pipe_pct_elapsed", "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"),
        ("l2_hit_pct", "lts__t_sector_hit_rate.pct"), ("sm_busy_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("issue_active_pct", "sm__inst_issued.avg.pct_of_peak_sustained_active"),
        ("achieved_warps_per_sm", "sm__warps_active.avg.per_cycle_active")]


def main():
    w = csv.writer(sys.stdout)
    w.writerow(["capture"] + [c for c, _ in COLS] + ["read_GBps"])
    for path in sys.argv[1:]:
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        idx = [hdr.index(m) if m in hdr else None for _, m in COLS]
        name = path.rsplit("/", 1)[-1].replace(".raw.csv", "")
        for r in rows[2:]:
            out = []
            for (c, _), i in zip(COLS, idx):
                v = r[i] if i is not None else ""
                if c == "kernel":
                    v = v.split("(")[0].replace("void ", "")
                elif v:
                    u = units[i]
                    f = float(v.replace(",", ""))
                    if c == "duration_us" and u in ("ns", "nsecond"):
                        f /= 1e3
                    if c.endswith("_MB") and u in ("byte", "B"):
                        f /= 1e6
                    if c.endswith("_MB") and u in ("Kbyte", "KB"):
                        f /= 1e3
                    v = f"{f:.0f}" if c in ("grid", "block", "regs", "occupancy_limit_blocks") else f"{f:.2f}"
                out.append(v)
            try:
                gbps = float(out[7]) / float(out[6]) * 1e3        # MB / us = TB/s ; x1e3 = GB/s
                out.append(f"{gbps:.0f}")
            except (ValueError, ZeroDivisionError):
                out.append("")
            w.writerow([name] + out)


if __name__ == "__main__":
    main()
