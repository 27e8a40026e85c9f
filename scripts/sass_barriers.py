"""Development aid: decode the scheduling control field of every SASS instruction of one
kernel (cuobjdump -sass output): stall count, write/read scoreboard set by the instruction
and the scoreboards it waits for.  Used to see which instructions wait on the scoreboard
of the global loads.
  cuobjdump -sass -fun <mangled> lib.so | python scripts/sass_barriers.py [pattern]
"""
import re
import sys

lines = sys.stdin.read().splitlines()
ins = []
i = 0
pat_ins = re.compile(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* 0x([0-9a-f]{16}) \*/")
pat_hi = re.compile(r"^\s+/\* 0x([0-9a-f]{16}) \*/")
while i < len(lines):
    m = pat_ins.match(lines[i])
    if m and i + 1 < len(lines):
        h = pat_hi.match(lines[i + 1])
        if h:
            hi = int(h.group(1), 16)
            ctrl = hi >> 41
            ins.append({
                "addr": m.group(1), "text": " ".join(m.group(2).split()),
                "stall": ctrl & 0xf, "yield": (ctrl >> 4) & 1,
                "wr": (ctrl >> 5) & 7, "rd": (ctrl >> 8) & 7, "wait": (ctrl >> 11) & 0x3f,
            })
            i += 2
            continue
    i += 1

show = sys.argv[1] if len(sys.argv) > 1 else None
for k, x in enumerate(ins):
    wr = "-" if x["wr"] == 7 else str(x["wr"])
    rd = "-" if x["rd"] == 7 else str(x["rd"])
    wait = "".join(str(b) for b in range(6) if (x["wait"] >> b) & 1) or "-"
    line = "%s  st%2d wr %s rd %s wait %-6s  %s" % (x["addr"], x["stall"], wr, rd, wait, x["text"])
    if show is None or re.search(show, x["text"]) or wait != "-":
        print(line)
