"""Dump one kernel of an `ncu --page source --csv` file: per SASS line executed count, samples,
avg threads.  usage: ncu_source_dump.py file.csv kernel_index [min_exec_share]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if 'Instructions Executed' in r]
n = int(sys.argv[2]); h = hi[n]; end = hi[n + 1] - 1 if n + 1 < len(hi) else len(rows)
H = rows[h]
iS = H.index('Source'); iE = H.index('Instructions Executed'); iSamp = H.index('# Samples'); iT = H.index('Avg. Threads Executed')
data = []
for r in rows[h + 1:end]:
    try: data.append((r[iS].strip(), int(r[iE]), int(r[iSamp]), float(r[iT])))
    except Exception: pass
tot = sum(d[1] for d in data); totS = sum(d[2] for d in data) or 1
print(rows[h - 1][1][:100], 'instr', tot, 'samples', totS)
for i, d in enumerate(data):
    print(f"{i:5d} {d[1]/tot*100:5.2f}% {d[2]/totS*100:5.2f}% {d[3]:4.1f}  {d[0][:90]}")
