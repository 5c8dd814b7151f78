"""Count the SASS mnemonics that prove the Blackwell-native paths, per kernel."""
import collections
import re
import subprocess
import sys

lib, out = sys.argv[1], sys.argv[2]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTC[A-Z]*MMA|UTCBAR|LDTM|UTMALDG|LDGSTS|SYNCS|HMMA|FFMA|RED|ATOMG|MULTIMEM[\w.]*|LDG[\w.]*\.SYS|STG[\w.]*\.SYS|LD\.E[\w.]*SYS[\w.]*|ST\.E[\w.]*SYS[\w.]*)\b")
cur, counts = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur:
        for t in pat.findall(line):
            counts[cur][t] += 1
demangled = {}
try:
    names = list(counts)
    d = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    demangled = dict(zip(names, d))
except Exception:
    pass
with open(out, "w") as fh:
    fh.write("SASS mnemonic counts per kernel (cuobjdump -sass, sm_100a)\n")
    fh.write("UTC*MMA = tcgen05.mma | LDTM = tcgen05.ld | UTCBAR = tcgen05.commit | UTMALDG = TMA load | LDGSTS = cp.async |\n"
             "SYNCS = mbarrier | MULTIMEM = NVLS multimem | *.SYS = system-scope (peer / NVLink) ld/st | FFMA = CUDA-core path\n\n")
    for k, c in counts.items():
        if c:
            fh.write(f"{demangled.get(k, k)[:150]}\n    " + ", ".join(f"{m} x{n}" for m, n in sorted(c.items())) + "\n")
print(sum(1 for c in counts.values() if c), "kernels ->", out)
