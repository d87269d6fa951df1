"""Static SASS instruction counts of one kernel per CUDA source file / line (nvdisasm -g over the cubin inside an object
file built with -lineinfo).  A cheap CPU-side proxy for comparing two versions of straight-line device code before any
GPU time is spent:  python tools/sass_lines.py build/se_ns.o ns_stream_kernelILi16EsLi0 [source-file-for-per-line-list]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

obj, fun = sys.argv[1], sys.argv[2]
d = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
cur = None
inside = False
cnt = collections.Counter()
byfile = collections.Counter()
ops = collections.defaultdict(collections.Counter)
for l in out.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        inside = fun in m.group(1)
        continue
    if l.lstrip().startswith(".section"):
        inside = False
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
    if m and cur:
        cnt[cur] += 1
        byfile[cur[0]] += 1
        ops[cur[0]][m.group(2).split(".")[0]] += 1
print("total", sum(cnt.values()))
for f, n in byfile.most_common():
    print(f, n, dict(ops[f].most_common(10)))
if len(sys.argv) > 3:
    for (f, ln), n in sorted(cnt.items()):
        if f == sys.argv[3]:
            print(f, ln, n)
