"""Development aid: prints the kernel durations of the last step in an ncu launch list
(ncu --metrics gpu__time_duration.sum --csv --log-file X ...).   python tools/launch_times.py X.csv"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
seq = [(r[ki][:44], float(r[vi].replace(',', '')) * (1e-3 if r[ui] == 'ns' else 1)) for r in rows[1:] if r[ki] != 'Kernel Name']
idx = [i for i, (k, _) in enumerate(seq) if 'k_project' in k]
for k, v in seq[idx[-2]:idx[-1]]:
    print("%-46s %9.1f us" % (k, v))
