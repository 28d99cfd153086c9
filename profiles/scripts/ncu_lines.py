"""Per-source-line instruction / stall-sample totals of every kernel in an .ncu-rep captured with
--import-source on (source page, cuda+sass view). Usage: ncu_lines.py report.ncu-rep [units] [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'], capture_output=True, text=True).stdout
kern = []; fpath = None
for r in csv.reader(raw.splitlines()):
    if r and r[0] == 'File Path': fpath = r[1].split('/')[-1]
    elif r and r[0] == 'Function Name':
        if not kern or kern[-1][0] != r[1]: kern.append((r[1], []))
    elif len(r) > 10 and r[0] not in ('', 'Line No') and kern:
        try: kern[-1][1].append((int(r[7]), fpath, int(r[0]), r[1][:100], int(r[6])))
        except ValueError: pass
for fn, out in kern:
    tot = sum(o[0] for o in out); ts = max(1, sum(o[4] for o in out))
    print('=====', fn[:70], 'warp-instr/unit %.1f' % (tot / units), 'samples', ts)
    for v, f, l, src, sm in sorted(out, reverse=True)[:top]:
        print('%5.1f%% (%7.1f/unit) smp %5.1f%% %s:%d %s' % (100 * v / tot, v / units, 100 * sm / ts, f, l, src))
