"""Per-source-line totals (instructions executed, stall samples) of an .ncu-rep captured with --import-source on.
   python tools/ncu_lines.py file.ncu-rep [top_n] [kernel substring]   (first launch of the matching kernel)"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
want = sys.argv[3] if len(sys.argv) > 3 else ''
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, out, fname, func, seen_main = None, [], '', '', {}
def num(x):
    try: return int(x)
    except ValueError: return 0
for r in rows:
    if r and r[0] == 'File Path': fname = r[1].split('/')[-1]
    if r and r[0] == 'Function Name':
        func = r[1]
        key = (func, fname)
        seen_main[key] = seen_main.get(key, 0) + 1
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr and r and r[0].isdigit() and want in func and seen_main.get((func, fname), 0) == 1:
        out.append((fname, int(r[0]), r[1].strip()[:100], num(r[hdr.index('Instructions Executed')]), num(r[hdr.index('# Samples')])))
tot = sum(o[3] for o in out) or 1; ts = sum(o[4] for o in out) or 1
print('warp instructions %d, samples %d' % (tot, ts))
for o in sorted(out, key=lambda o: -o[4])[:top]:
    print('%s:%4d %5.1f%% inst %5.1f%% samp  %s' % (o[0], o[1], 100 * o[3] / tot, 100 * o[4] / ts, o[2]))
