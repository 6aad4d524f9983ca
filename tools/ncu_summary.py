#!/usr/bin/env python
"""Turn an .ncu-rep (ncu --set full) into the markdown table kept under profiles/.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep [title] > profiles/rNN_x.md
Needs the `ncu` CLI (no GPU): it only reads the report."""
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(r"^(Kernel Name|gpu__time_duration\.sum|dram__bytes_(read|write)\.sum(\.per_second)?|dram__sectors_(read|write)\.sum|"
                  r"gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|lts__t_sector_hit_rate\.pct|lts__t_sectors_srcunit_tex_op_read\.sum|"
                  r"lts__t_sectors_op_(atom|red)\.sum|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|"
                  r"l1tex__t_(sectors|requests)_pipe_lsu_mem_global_op_(ld|st|red|atom)\.sum|launch__(grid_size|block_size|registers_per_thread|occupancy_limit_\w+)|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|sm__cycles_elapsed\.avg\.per_second|"
                  r"smsp__inst_executed\.sum|smsp__issue_active\.avg\.pct_of_peak_sustained_active|"
                  r"smsp__average_warps_issue_stalled_(long_scoreboard|lg_throttle|short_scoreboard|wait|math_pipe_throttle|not_selected|mio_throttle)_per_issue_active\.ratio)$")


def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2:]
    print("# %s\n" % title)
    print("| metric | unit | value(s) |\n|---|---|---|")
    for i, name in enumerate(hdr):
        if KEEP.match(name):
            print("| %s | %s | %s |" % (name, units[i], ", ".join(v[i] for v in vals)))


if __name__ == "__main__":
    main()
