#!/usr/bin/env python
"""Attribute an ncu capture's per-SASS-instruction counters to CUDA source lines.

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [--kernel REGEX] [--top N] [--lib PATH]

ncu's `--page source --csv` lists SASS instructions with executed-instruction and stall-sample counts
but (in CSV form) not the source line; `nvdisasm -gi` on the cubin embedded in the library gives the
line (and inlining chain) of every SASS instruction of the same build.  Both list a kernel's
instructions in address order, so they are joined by position.  Output: per kernel, the share of
executed warp-instructions and of stall samples per source line / per outermost inlined call site.
The library must be the same build the capture ran (compile with -lineinfo).
"""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_lines(lib):
    """{mangled kernel name: [(file, line, outer_file, outer_line, text), ...]} in address order."""
    tmp = tempfile.mkdtemp(prefix="xelf")
    subprocess.check_call(["cuobjdump", "-xelf", "all", lib], cwd=tmp, stdout=subprocess.DEVNULL)
    out = {}
    for f in sorted(os.listdir(tmp)):
        p = os.path.join(tmp, f)
        if not f.endswith(".cubin") or os.path.getsize(p) == 0:
            continue
        txt = subprocess.run(["nvdisasm", "-c", "-gi", p], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
        cur, loc, grp, fresh = None, ("?", 0, "?", 0, "?", 0), [], True
        for ln in txt.splitlines():
            m = re.match(r"^\.text\.(\S+):", ln)
            if m:
                cur = m.group(1)
                out[cur] = []
                continue
            if cur is None:
                continue
            # the inlining chain of the next instruction: consecutive "//## File" lines, innermost first
            m = re.match(r'^\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                if fresh:
                    grp, fresh = [], False
                grp.append((os.path.basename(m.group(1)), int(m.group(2))))
                continue
            m = re.match(r"^\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
            if m:
                if grp:
                    mid = grp[-2] if len(grp) >= 2 else grp[-1]
                    loc = grp[0] + mid + grp[-1]        # innermost, one level below the kernel body, kernel body
                fresh = True
                out[cur].append(loc + (m.group(1).strip(),))
    return out


def ncu_source(rep):
    """[(kernel display name, [row dict, ...])] from `ncu --page source --csv`."""
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    kernels, hdr, cur = [], None, None
    for row in csv.reader(io.StringIO(txt)):
        if not row:
            continue
        if row[0] == "Kernel Name":
            cur = (row[1], [])
            kernels.append(cur)
            hdr = None
        elif row[0] == "Address":
            hdr = row
        elif hdr is not None and cur is not None:
            cur[1].append(dict(zip(hdr, row)))
    return kernels


def demangle(names):
    p = subprocess.run(["cu++filt"] + list(names), stdout=subprocess.PIPE, text=True)
    return dict(zip(names, p.stdout.splitlines()))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--lib", default=os.path.join(ROOT, "cvxcompress_b200", "lib", "libcvxcompress.so"))
    ap.add_argument("--kernel", default=".*")
    ap.add_argument("--top", type=int, default=25)
    ap.add_argument("--focus", default=None, help="file:line of the kernel body: stall reasons and opcode mix of that call site only")
    args = ap.parse_args()
    sass = sass_lines(args.lib)
    dm = demangle(list(sass.keys()))
    norm = lambda s: re.sub(r"\((int|bool|long)\)", "", s).replace("void ", "").replace(" ", "")
    by_dm = {norm(v): k for k, v in dm.items()}
    seen = set()
    for name, rows in ncu_source(args.report):
        if not re.search(args.kernel, name) or name in seen:
            continue
        seen.add(name)
        key = by_dm.get(norm(name))
        print("=" * 100)
        print(name, "--", len(rows), "SASS instructions in the capture")
        if key is None or len(sass[key]) != len(rows):
            print("  cannot join with the library's SASS (%s instructions there): rebuild mismatch?" %
                  (len(sass[key]) if key else "kernel not found;"))
            continue
        tot_i = sum(float(r["Instructions Executed"] or 0) for r in rows) or 1.0
        tot_s = sum(float(r["# Samples"] or 0) for r in rows) or 1.0
        inner = collections.defaultdict(lambda: [0.0, 0.0, 0.0])
        mid = collections.defaultdict(lambda: [0.0, 0.0, 0.0])
        outer = collections.defaultdict(lambda: [0.0, 0.0, 0.0])
        stalls = collections.Counter()
        opc = collections.Counter()
        if args.focus:
            ff, fl = args.focus.rsplit(":", 1)
            pairs = [(r, s) for r, s in zip(rows, sass[key]) if s[4].endswith(ff) and s[5] == int(fl)]
            fi = sum(float(r["Instructions Executed"] or 0) for r, _ in pairs)
            fs = sum(float(r["# Samples"] or 0) for r, _ in pairs) or 1.0
            st, oc, sub = collections.Counter(), collections.Counter(), collections.defaultdict(lambda: [0.0, 0.0])
            for r, s in pairs:
                ie = float(r["Instructions Executed"] or 0)
                oc[s[6].split()[0] if not s[6].startswith("@") else s[6].split()[1]] += ie
                sub[(s[2], s[3])][0] += ie; sub[(s[2], s[3])][1] += float(r["# Samples"] or 0)
                for c, v in r.items():
                    if c.startswith("stall_") and "Not Issued" not in c and v:
                        st[c] += float(v)
            print("  focus %s: %d SASS, inst %.2f%% of kernel, samples %.2f%%" % (args.focus, len(pairs), 100 * fi / tot_i, 100 * fs / tot_s))
            print("    stalls: " + ", ".join("%s %.1f%%" % (k[6:], 100 * v / fs) for k, v in st.most_common(10)))
            print("    opcodes: " + ", ".join("%s %.1f%%" % (k, 100 * v / (fi or 1)) for k, v in oc.most_common(16)))
            for k, v in sorted(sub.items(), key=lambda kv: -kv[1][1])[:args.top]:
                print("     %6.2f  %6.2f   %s:%d" % (100 * v[0] / (fi or 1), 100 * v[1] / fs, k[0], k[1]))
            continue
        for r, s in zip(rows, sass[key]):
            ie, ns, te = float(r["Instructions Executed"] or 0), float(r["# Samples"] or 0), float(r["Thread Instructions Executed"] or 0)
            for d, k in ((inner, (s[0], s[1])), (mid, (s[2], s[3])), (outer, (s[4], s[5]))):
                d[k][0] += ie; d[k][1] += ns; d[k][2] += te
            opc[s[6].split()[0] if not s[6].startswith("@") else s[6].split()[1]] += ie
            for c, v in r.items():
                if c.startswith("stall_") and "Not Issued" not in c and v:
                    stalls[c] += float(v)
        print("  warp-instructions executed: %.4g   stall samples: %.4g" % (tot_i, tot_s))
        print("  stall reasons: " + ", ".join("%s %.1f%%" % (k[6:], 100 * v / tot_s) for k, v in stalls.most_common(8)))
        print("  opcodes: " + ", ".join("%s %.1f%%" % (k, 100 * v / tot_i) for k, v in opc.most_common(14)))
        for title, d in (("by line of the kernel body", outer), ("one inlining level below the kernel body", mid),
                         ("by innermost source line", inner)):
            print("  -- %s: inst%%  samples%%  avg-active-lanes  file:line" % title)
            for k, v in sorted(d.items(), key=lambda kv: -kv[1][1])[:args.top]:
                print("     %6.2f  %6.2f  %5.1f   %s:%d" % (100 * v[0] / tot_i, 100 * v[1] / tot_s, v[2] / v[0] if v[0] else 0, k[0], k[1]))


if __name__ == "__main__":
    main()
