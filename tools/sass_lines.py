"""SASS instructions per source line of one kernel, without a GPU: python tools/sass_lines.py <obj.cu.o> <kernel substring> [file substring]
(nvdisasm --print-line-info on the cubin inside the object; inlined code is attributed to the innermost line)."""
import collections
import os
import re
import subprocess
import sys
import tempfile

obj, kern = sys.argv[1], sys.argv[2]
fsub = sys.argv[3] if len(sys.argv) > 3 else ""
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
cnt = collections.Counter()
inside = False
cur = None
for line in txt.splitlines():
    if line.startswith("//---") and ".text." in line:
        inside = kern in line
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", line):
        cnt[cur] += 1
tot = sum(cnt.values())
print("total", tot)
for (f, l), n in sorted(cnt.items()):
    if fsub in f:
        print("%s:%d %d" % (f, l, n))
