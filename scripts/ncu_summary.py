"""Summarise an .ncu-rep (or its `--page raw --csv` export) into a per-launch table:
    python scripts/ncu_summary.py file.ncu-rep | file_raw.csv"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = (open(rep).read() if rep.endswith(".csv") else
       subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64%"),
        ("lts__t_sector_hit_rate.pct", "l2hit%"), ("l1tex__t_sector_hit_rate.pct", "l1hit%"),
        ("launch__registers_per_thread", "regs"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
        ("lts__t_requests_srcunit_tex.sum", "l2_requests")]
ki = hdr.index("Kernel Name")
print("| kernel | " + " | ".join(n for _, n in want) + " |")
print("|---|" + "---|" * len(want))
for r in rows[2:]:
    cells = []
    for m, _ in want:
        if m in hdr:
            i = hdr.index(m)
            v = r[i]
            try:
                v = f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
            cells.append(f"{v} {units[i]}".strip())
        else:
            cells.append("n/a")
    name = r[ki].split("(")[0].replace("void ", "").replace("efgp::", "")
    print(f"| {name} | " + " | ".join(cells) + " |")
