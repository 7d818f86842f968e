import csv, sys, subprocess
rep, out = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines())); h = rows[0]
cols = ['Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_issued.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','launch__block_size']
with open(out, 'w', newline='') as f:
    w = csv.writer(f)
    for r in rows: w.writerow([r[h.index(c)] if c in h else '' for c in cols])
print(open(out).read())
