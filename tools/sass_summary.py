#!/usr/bin/env python
"""SASS evidence for the Blackwell-specific claims of DESIGN.md, reproducible from the repo (no GPU needed):
per kernel of cvxcompress_b200/lib/libcvxcompress.so, the count of the instructions that prove TMA loads (UTMALDG), bulk copies (UBLKCP),
mbarrier waits (SYNCS), packed fp32 arithmetic (FFMA2 / FADD2 / FMUL2), cp.async (LDGSTS), named barriers (BAR),
warp votes and shuffles; plus registers and spills from the ptxas log.   python tools/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "cvxcompress_b200", "lib", "libcvxcompress.so")
ops = ["UTMALDG", "UBLKCP", "SYNCS", "FFMA2", "FADD2", "FMUL2", "FFMA", "FADD", "LDGSTS", "LDS", "STS", "LDG", "STG", "BAR", "VOTE", "SHFL", "ATOMG", "REDUX"]
txt = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True).stdout
cur, table = None, collections.OrderedDict()
for ln in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("void ", "")
        table[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", ln)
    if m and cur:
        table[cur][m.group(1)] += 1
        table[cur]["_total"] += 1
print("library:", os.path.relpath(lib, ROOT))
print("%-58s %7s " % ("kernel", "SASS") + " ".join("%7s" % o for o in ops))
for k, c in table.items():
    if c["_total"] < 200:
        continue
    print("%-58s %7d " % (k[:58], c["_total"]) + " ".join("%7d" % c[o] for o in ops))
log = os.path.join(os.path.dirname(lib), "obj", "ptxas_kernels.log")
if os.path.exists(log):
    print("\nregisters / spills (ptxas -v):")
    name = None
    for ln in open(log):
        m = re.search(r"Compiling entry function '(\S+)'", ln)
        if m:
            name = re.sub(r"\(.*", "", subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()).replace("void ", "")
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", ln)
        if m and name:
            spill = (m.group(2), m.group(3))
        m = re.search(r"Used (\d+) registers", ln)
        if m and name:
            print("  %-70s %3s registers, spill stores/loads %s/%s bytes" % (name[:70], m.group(1), spill[0], spill[1]))
            name = None
