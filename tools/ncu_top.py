"""Prints the hottest SASS instructions (by warp-stall samples) of every kernel in an `ncu --page source --csv` dump."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = None
data = []
def flush():
    if not data: return
    tot = sum(d[0] for d in data) or 1
    print('total samples', tot)
    for s, src, ex, k in sorted(data, reverse=True)[:n]:
        print(f'{s:7d} {100*s/tot:5.1f}%  #{k:5d} exec={ex:>9}  {src[:120]}')
for r in rows:
    if r and r[0] == 'Kernel Name':
        flush(); data = []; print('\n====', r[1][:120]); continue
    if r and r[0] == 'Address':
        hdr = r; i_src = hdr.index('Source'); i_s = hdr.index('# Samples'); i_ex = hdr.index('Instructions Executed'); continue
    if hdr and len(r) > i_s and r[i_s].isdigit():
        data.append((int(r[i_s]), r[i_src].strip(), r[i_ex], len(data)))
flush()
