"""Static SASS opcode histogram of six k_voxel_update_tiled instantiations (cuobjdump -sass of the built object).
usage: python profiles/scripts/sass_histogram.py > profiles/r02_update_sass_histogram.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
OBJ = os.path.join(ROOT, "voxel_game_b200", "lib", "k_update.o")
sass = subprocess.run(["cuobjdump", "-sass", OBJ], capture_output=True, text=True, check=True).stdout
funcs, cur = {}, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        funcs[cur].append(m.group(1))
print("Static SASS opcode histogram of k_voxel_update_tiled<PHASE, ELIDE, LINK> in libvoxelb200.so (cuobjdump -sass, round 2, final build;")
print("profiles/scripts/sass_histogram.py). The staging is LDGSTS (cp.async, 16-byte + 4-byte forms); no UBLKCP / UTMALDG: the copies are")
print("thread-private vectors, see DESIGN.md 4.1.\n")
for phase, elide, link in ((7, 0, 1), (4, 0, 1), (0, 0, 0), (7, 1, 0), (4, 1, 0), (0, 1, 0)):
    name = "_ZN3vxb20k_voxel_update_tiledILi%dELb%dELb%dEEE" % (phase, elide, link)
    ops = next(v for k, v in funcs.items() if k.startswith(name))
    h = collections.Counter(ops)
    print("<%d, %s, %s>  %d instructions" % (phase, "elide" if elide else "store", "link" if link else "no link", len(ops)))
    print("  " + ", ".join("%s %d" % kv for kv in h.most_common(28)))
    print("  " + "  ".join("%s %d" % (k, h.get(k, 0)) for k in ("LDGSTS", "LDS", "STS", "SHFL", "BAR", "VOTE", "UBLKCP", "UTMALDG")) + "\n")
