#!/usr/bin/env python
"""Decode the scheduling control bits of sm_100 SASS (stall count, write/read barrier index, wait mask) for one kernel:
   python tools/sass_ctrl.py <lib.so> <mangled-name-substring> [regex-filter]
Used to see which scoreboard barrier the record loads of the VBD streaming kernel are assigned and where it is waited on."""
import re, subprocess, sys
lib, name = sys.argv[1], sys.argv[2]
flt = re.compile(sys.argv[3]) if len(sys.argv) > 3 else None
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
blocks = out.split('Function : ')
for b in blocks[1:]:
    fn = b.split('\n', 1)[0]
    if name not in fn:
        continue
    print('==', fn[:120])
    lines = b.split('\n')
    i = 0
    n = 0
    while i < len(lines):
        m = re.match(r'\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/', lines[i])
        if m and i + 1 < len(lines):
            m2 = re.match(r'\s+/\* (0x[0-9a-f]{16}) \*/', lines[i + 1])
            if m2:
                hi = int(m2.group(1), 16)
                ctrl = hi >> 41
                stall = ctrl & 0xf
                wr = (ctrl >> 5) & 7
                rd = (ctrl >> 8) & 7
                wait = (ctrl >> 11) & 0x3f
                n += 1
                txt = m.group(2)
                if flt is None or flt.search(txt) or wait:
                    if flt is None or flt.search(txt):
                        print(f'{n:5d} {m.group(1)} st{stall:2d} wr{wr if wr != 7 else "-"} rd{rd if rd != 7 else "-"} wait{wait:06b}  {txt[:90]}')
                i += 2
                continue
        i += 1
    break
