"""Per-SASS-instruction view of an .ncu-rep (source page): shared-memory bank conflicts, excessive L2 sectors and the
opcode histogram (executed warp instructions per opcode). Usage: ncu_sass_hot.py report.ncu-rep [kernel-substring] [units]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]; pick = sys.argv[2] if len(sys.argv) > 2 else ''; units = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
kern = None; hdr = None; data = collections.OrderedDict()
for r in rows:
    if r and r[0] == 'Kernel Name': kern = r[1]; data[kern] = []; hdr = None
    elif r and r[0] == 'Address': hdr = {h: i for i, h in enumerate(r)}
    elif kern and hdr and len(r) > 20: data[kern].append(r)
for k, rs in data.items():
    if pick not in k: continue
    print('=====', k[:100])
    ops = collections.Counter(); tot = 0
    for r in rs:
        n = int(r[hdr['Instructions Executed']]); tot += n
        src = r[hdr['Source']].split()
        op = src[1] if src and src[0].startswith('@') else (src[0] if src else '?')
        ops[op.split('.')[0]] += n
    print('executed warp instructions: %d (%.1f per unit); by opcode:' % (tot, tot / units))
    print('  ' + ', '.join('%s %.1f%%' % (o, 100.0 * n / tot) for o, n in ops.most_common(22)))
    for col, title in (('L1 Conflicts Shared N-Way', 'shared bank conflicts'), ('L2 Theoretical Sectors Global Excessive', 'excessive L2 sectors'), ('stall_mio', 'mio stall samples'), ('stall_short_sb', 'short-scoreboard samples'), ('stall_barrier', 'barrier samples')):
        if col not in hdr: continue
        best = sorted(((int(float(r[hdr[col]] or 0)), r[hdr['Source']].strip(), int(r[hdr['Instructions Executed']])) for r in rs), reverse=True)[:8]
        if best and best[0][0] > 0:
            print('  top', title)
            for v, s, n in best:
                if v > 0: print('    %12d  (exec %9d)  %s' % (v, n, s[:110]))
