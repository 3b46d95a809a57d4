"""Instruction count per source line and per function of a cubin (nvdisasm -g): where the code size of a kernel goes.
usage: python tools/sass_lines.py file.cubin [top]"""
import collections
import re
import subprocess
import sys

txt = subprocess.run(["nvdisasm", "-g", sys.argv[1]], capture_output=True, text=True).stdout.split("\n")
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None
func = None
cnt = collections.Counter()
fcnt = collections.Counter()
for l in txt:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"^(\$?[_A-Za-z]\S*):", l)
    if m and not m.group(1).startswith(".L"):
        func = m.group(1)
    if re.match(r"^\s+/\*[0-9a-f]{4,}\*/\s+[A-Z@]", l) and l.rstrip().endswith(";"):
        cnt[cur] += 1
        fcnt[func] += 1
print("total instructions", sum(cnt.values()))
for k, v in fcnt.most_common(20):
    print(f"  {v:6d} ({v * 16 / 1024:6.1f} KB)  {k}")
for k, v in cnt.most_common(top):
    print(f"  {v:6d}  {k}")
