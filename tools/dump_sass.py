#!/usr/bin/env python
"""Write one SASS listing per kernel of libgk_cuda.so into profiles/sass/ plus an index with the
instruction mix (what proves which memory / synchronisation instructions a kernel really uses).

    python tools/dump_sass.py            # needs cuobjdump + c++filt on PATH; no GPU
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "gaussiank_sgd_b200", "ops", "libgk_cuda.so")
OUT = os.path.join(ROOT, "profiles", "sass")
WATCH = ["LDG", "STG", "LDGSTS", "LDS", "STS", "ATOMG", "RED", "ATOMS", "BAR", "MEMBAR", "FENCE", "ERRBAR", "CCTL",
         "SHFL", "VOTE", "MATCH", "NANOSLEEP", "MUFU", "DFMA", "DMUL", "DADD", "FFMA", "IMAD", "BRA",
         "UTCHMMA", "UTCBAR", "UTCATOMSWS", "UTMALDG", "UTMACCTL", "UBLKCP", "LDTM", "SYNCS", "ELECT"]


def demangle(name):
    try:
        return subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
    except OSError:
        return name


def short(dem):
    base = dem.replace("(anonymous namespace)::", "").split("(")[0].replace("void ", "")
    return re.sub(r"[^A-Za-z0-9_]+", "_", base).strip("_")


def main():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        if f.endswith(".sass") or f == "INDEX.txt":
            os.remove(os.path.join(OUT, f))
    parts = re.split(r"\n\s*Function : ", txt)
    index = []
    for part in parts[1:]:
        name, _, body = part.partition("\n")
        dem = demangle(name.strip())
        lines = []
        for ln in body.splitlines():
            s = ln.rstrip()
            if not s.strip() or s.lstrip().startswith("//##"):      # drop the -lineinfo file/line markers
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4})\*/\s+(.*?)\s*/\* 0x[0-9a-f]+ \*/", s)
            if m:
                lines.append("/*%s*/  %s" % (m.group(1), m.group(2).rstrip(" ;") + " ;"))
            elif re.match(r"\s*/\* 0x[0-9a-f]+ \*/", s):             # second encoding word
                continue
            elif s.lstrip().startswith("."):
                lines.append(s.strip())
        mix = collections.Counter()
        for ln in lines:
            m = re.match(r"/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
            if m:
                op = m.group(1)
                for w in WATCH:
                    if op == w or op.startswith(w + "."):
                        mix[w] += 1
        fn = short(dem) + ".sass"
        with open(os.path.join(OUT, fn), "w") as f:
            f.write("// %s\n// %s\n// sm_100a, %d instructions\n" % (dem, name.strip(), sum(1 for x in lines if x.startswith("/*"))))
            f.write("\n".join(lines) + "\n")
        index.append((fn, dem, sum(1 for x in lines if x.startswith("/*")), mix))
    with open(os.path.join(OUT, "INDEX.txt"), "w") as f:
        f.write("SASS listings of gaussiank_sgd_b200/ops/libgk_cuda.so (nvcc 12.9, -gencode arch=compute_100a,code=sm_100a -O3).\n"
                "Per kernel: file, instruction count, and how many instructions of each watched class it contains.\n\n")
        for fn, dem, n, mix in sorted(index):
            f.write("%s\n    %s\n    %d instructions; %s\n" % (fn, dem, n, ", ".join("%s %d" % (k, mix[k]) for k in WATCH if mix[k])))
    print("wrote %d listings to %s" % (len(index), OUT))


if __name__ == "__main__":
    sys.exit(main())
