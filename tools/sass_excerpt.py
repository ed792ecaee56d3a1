"""The Blackwell-native instructions of one kernel of the shipped library, for profiles/sass_<kernel>.txt.
   python tools/sass_excerpt.py k_net_trunk_rows > profiles/sass_k_net_trunk_rows.txt"""
import collections
import re
import subprocess
import sys

name = sys.argv[1]
sass = subprocess.run(["cuobjdump", "-sass", "alphazero_b200/libaz_b200.so"], capture_output=True, text=True).stdout
func, lines, inside = None, [], False
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        inside = name in m.group(1) and (func is None or func == m.group(1))
        if inside:
            func = m.group(1)
        continue
    if inside and re.search(r"/\*[0-9a-f]{4}\*/", l):
        lines.append(l)
KEYS = ("UTC", "LDTM", "STTM", "UBLKCP", "UCGABAR", "SYNCS", "ELECT", "R2UR", "MEMBAR", "FENCE", "REDUX", "REDG", "MAPA", "CCTL", "ACQBULK")
counts = collections.Counter()
picked = []
for l in lines:
    op = l.split("*/", 1)[1].strip().split(";")[0]
    mnem = next((t for t in op.split() if not t.startswith("@")), "")
    if any(k in mnem for k in KEYS):
        base = ".".join(mnem.split(".")[:2]) if mnem.startswith("SYNCS") or mnem.startswith("FENCE") else mnem.split(".")[0]
        counts[base] += 1
        if not mnem.startswith("R2UR"):
            picked.append(l)
print(f"# cuobjdump -sass alphazero_b200/libaz_b200.so, function {func}: {len(lines)} SASS instructions")
print("# Blackwell-native mnemonics (tcgen05.mma -> UTC*MMA, tcgen05.commit -> UTCBAR, tcgen05.ld -> LDTM, tcgen05.alloc -> UTCATOMSWS,")
print("# mbarrier -> SYNCS, fence.proxy.async -> FENCE.VIEW.ASYNC); counts, then the lines in program order")
print("# " + ", ".join(f"{k}: {v}" for k, v in counts.most_common()))
for l in picked:
    print(l.rstrip())
