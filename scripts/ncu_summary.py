#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into a small text file for profiles/.

  python scripts/ncu_summary.py gpurun_out/prof_write.ncu-rep > profiles/rNN_<kernel>.txt
"""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "lts__t_bytes.sum"]
STALLS = ["stall_lg", "stall_long_sb", "stall_math", "stall_mio", "stall_short_sb", "stall_wait",
          "stall_not_selected", "stall_selected", "stall_dispatch", "stall_barrier", "stall_branch_resolving",
          "stall_no_inst", "stall_drain", "stall_membar", "stall_sleep", "stall_tex", "stall_misc"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows = page(rep, "raw")
    h, units = rows[0], rows[1]
    print("# ncu --set full --clock-control none --import-source on   (%s)" % rep)
    for k, r in enumerate(rows[2:]):
        print("\n## launch %d: %s" % (k, r[h.index("Kernel Name")][:100]))
        for w in WANT:
            if w in h:
                print("%-62s %s %s" % (w, r[h.index(w)], units[h.index(w)]))
        if "dram__bytes_read.sum" in h:
            def val(name):
                v, u = float(r[h.index(name)].replace(",", "")), units[h.index(name)]
                return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
            tr = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
            us = float(r[h.index("gpu__time_duration.sum")].replace(",", ""))
            scale = {"us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(units[h.index("gpu__time_duration.sum")], 1e-6)
            print("%-62s %.0f bytes  -> %.1f GB/s under ncu" % ("traffic (dram read+write)", tr, tr / (us * scale) / 1e9))
    src = page(rep, "source")
    if len(src) > 2:
        h2 = src[1]
        n = 0
        tot = {s: 0 for s in STALLS}
        samples = 0
        for r in src[2:]:
            if r and r[0] == "Kernel Name":
                break
            if len(r) < len(h2):
                continue
            n += 1
            try:
                samples += int(r[h2.index("# Samples")])
            except ValueError:
                continue
            for s in STALLS:
                if s in h2:
                    try:
                        tot[s] += int(r[h2.index(s)])
                    except ValueError:
                        pass
        print("\n## warp-stall sampling, launch 0 (%d SASS instructions, %d samples)" % (n, samples))
        for s, v in sorted(tot.items(), key=lambda kv: -kv[1]):
            if v:
                print("%-28s %6d  %5.1f%%" % (s, v, 100.0 * v / max(samples, 1)))


if __name__ == "__main__":
    main()
