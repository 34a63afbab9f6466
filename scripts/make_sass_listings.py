#!/usr/bin/env python
"""SASS evidence for every CUDA source (no GPU needed): for each build/<name>.o
  profiles/sass/<name>.sass.gz        the complete `cuobjdump -sass` listing, encodings stripped
  profiles/sass/<name>.kernels.txt    per kernel: instruction count and the mnemonics that matter
                                      (UTC*MMA / LDTM / STTM / UBLKCP / UTMA* / REDG / ATOMG / LDG.E.128 ...)
  profiles/sass/<kernel>.sass         plain-text listing of the headline kernels
Run after build_native.py."""
import collections
import gzip
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "sass")
PLAIN = ["k_mf_tiles_tma", "k_mf_sgdILi32ELi1ELb0ELi1ELi8ELb0", "k_lemp_topk_packed", "k_bucket_scatter",
         "k_pa_binary", "k_neg_draw", "k_lemp_topkILb1"]
KEY = re.compile(r"^(UTC\w*MMA|UTCBAR|LDTM|STTM|UBLKCP|UTMA\w+|SYNCS|REDG|RED|ATOMG|ATOM|ATOMS|LDG|STG|LDS|STS|HMMA|FFMA|SHFL|BAR|MEMBAR|"
                 r"CCTL|LDGSTS|REDUX|ELECT)")


def demangle(names):
    try:
        r = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True)
        return r.stdout.strip().split("\n")
    except Exception:
        return names


def main():
    os.makedirs(OUT, exist_ok=True)
    build = os.path.join(ROOT, "build")
    for obj in sorted(f for f in os.listdir(build) if f.endswith(".o")):
        name = obj[:-2]
        txt = subprocess.run(["cuobjdump", "-sass", os.path.join(build, obj)], capture_output=True, text=True).stdout
        lines = []
        for ln in txt.splitlines():
            if re.match(r"^\s*/\* 0x[0-9a-f]{16} \*/\s*$", ln):
                continue                                    # second encoding word
            lines.append(re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", ln))
        with gzip.open(os.path.join(OUT, name + ".sass.gz"), "wt") as f:
            f.write("\n".join(lines) + "\n")
        kernels, cur = collections.OrderedDict(), None
        for ln in lines:
            m = re.search(r"Function : (\S+)", ln)
            if m:
                cur = m.group(1)
                kernels[cur] = []
                continue
            m = re.match(r"^\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]*)", ln)
            if m and cur is not None:
                kernels[cur].append((m.group(1), ln))
        pretty = demangle(list(kernels))
        with open(os.path.join(OUT, name + ".kernels.txt"), "w") as f:
            for (mangled, ins), nice in zip(kernels.items(), pretty):
                hist = collections.Counter(op for op, _ in ins)
                key = {k: v for k, v in hist.items() if KEY.match(k)}
                wide = sum(v for k, v in hist.items() if k.startswith(("LDG", "REDG", "ATOMG", "STG")) and ".128" in k or "F32x4" in k)
                f.write(f"{nice}\n  {len(ins)} instructions; 128-bit global accesses: {wide}\n")
                f.write("  " + ", ".join(f"{k} x{v}" for k, v in sorted(key.items())) + "\n\n")
                for want in PLAIN:
                    if want in mangled:
                        base = re.search(r"(k_\w+)", nice).group(1)
                        targs = re.search(r"<([^>]*)>", nice)
                        short = base + ("_" + re.sub(r"[^A-Za-z0-9]+", "_", targs.group(1)).strip("_") if targs else "")
                        with open(os.path.join(OUT, f"{name}.{short}.sass"), "w") as g:
                            g.write(f"// {nice}\n// from build/{obj} (nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo)\n")
                            g.write("\n".join(l for _, l in ins) + "\n")
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
