"""Summarise an .ncu-rep (raw page) into a markdown table: python scripts/ncu_summary.py rep [rep...]"""
import csv, subprocess, sys
WANT = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
        ("launch__registers_per_thread", "regs"), ("launch__waves_per_multiprocessor", "waves")]
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    print(f"\n### {rep.split('/')[-1]}\n")
    print("| kernel | grid | " + " | ".join(n for _, n in WANT) + " |")
    print("|---|---|" + "---|" * len(WANT))
    for r in rows[2:]:
        d = dict(zip(hdr, r)); u = dict(zip(hdr, units))
        cells = []
        for k, _ in WANT:
            v = d.get(k, "")
            try:
                v = f"{float(v):.2f}"
            except ValueError:
                pass
            cells.append(f"{v} {u.get(k, '') if 'bytes' in k else ''}".strip())
        print(f"| `{d['Kernel Name'][:48]}` | {d['Grid Size']} | " + " | ".join(cells) + " |")
