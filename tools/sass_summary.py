"""Per-kernel SASS mnemonic evidence (tcgen05 / TMA / tensor-core ops) from the built library -> profiles/sass_summary.txt"""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "deeperspeed_b200", "lib", "libdsb200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTC[A-Z0-9]*MMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|UBLKCP|UBLKRED|LDGMC|HMMA|SYNCS|LDGSTS|RED|ATOM)\b")
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur:
        for mm in pat.findall(line):
            counts[cur][mm] += 1
with open(os.path.join(root, "profiles", "sass_summary.txt"), "w") as f:
    f.write("SASS mnemonic counts per kernel (cuobjdump -sass libdsb200.so, sm_100a)\n")
    for k, c in counts.items():
        dem = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()[:110]
        f.write(f"{dem}\n    {dict(c)}\n")
print("kernels:", len(counts))
