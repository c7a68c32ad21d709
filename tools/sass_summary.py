"""SASS evidence for every kernel in libdml_b200.so (cuobjdump must be on PATH or under /usr/local/cuda/bin).

    python tools/sass_summary.py [--out-dir profiles] [--tag r02]

Writes  <out-dir>/<tag>_sass_summary.csv      one row per kernel: instruction count and the counts of the mnemonics that
                                              prove (or rule out) the Blackwell tensor / TMA path (B200_PROFILING.md):
                                              UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tensor
                                              load/store, UBLKCP = bulk copy, SYNCS = mbarrier, HMMA = legacy mma.sync, ...
        <out-dir>/<tag>_sass_<kernel>.sass    for the kernels on the engine's path: the listing reduced to those
                                              instructions (with their addresses), so the MMA loops can be read at a glance.
"""
import argparse
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-machine-learning_b200", "libdml_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA",
        "LDGSTS", "FADD2", "HMNMX2", "ELECT", "ACQBULK", "LDG", "STG", "LDS", "STS"]
ENGINE_PATH = ["conv_tc2x_kernel", "conv_tc2_kernel", "conv_tc5_kernel", "conv_tc6_kernel", "conv_tc6x_kernel", "conv_chain_kernel", "conv_duo_kernel", "conv_tc6p_kernel",
               "preprocess_s2d_kernel", "maxpool3_bf16_kernel", "avgpool3_kernel", "gap_kernel", "fc_tf32_kernel", "softmax_top5_kernel"]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout.splitlines()
        return dict(zip(names, out))
    except Exception:
        return {n: n for n in names}


def short(name):
    m = re.search(r"(\w+)(<[^(]*>)?\(", name)
    base = m.group(1) if m else name
    targs = m.group(2) if m and m.group(2) else ""
    return base, re.sub(r"[^0-9A-Za-z]+", "_", targs).strip("_")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out-dir", default=os.path.join(ROOT, "profiles"))
    ap.add_argument("--tag", default="r02")
    args = ap.parse_args()
    cuobjdump = "cuobjdump" if subprocess.run(["which", "cuobjdump"], capture_output=True).returncode == 0 else "/usr/local/cuda/bin/cuobjdump"
    sass = subprocess.run([cuobjdump, "--dump-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur and re.match(r"\s*/\*[0-9a-f]{4,}\*/", line):
            funcs[cur].append(line.rstrip())
    names = demangle(list(funcs))
    rows = []
    for mangled, lines in funcs.items():
        base, targs = short(names[mangled])
        counts = {k: 0 for k in KEYS}
        for l in lines:
            ins = re.sub(r"/\*.*?\*/", "", l).strip()
            for k in KEYS:
                if re.search(r"(^|[\s@!P0-9])" + k + r"(\.|\s|$)", ins):
                    counts[k] += 1
        rows.append((base, targs, len(lines), counts))
        if base in ENGINE_PATH:
            fn = os.path.join(args.out_dir, f"{args.tag}_sass_{base}{('_' + targs) if targs else ''}.sass")
            keep = [l for l in lines if re.search(r"UTC|LDTM|STTM|UTMA|UBLKCP|SYNCS|HMMA|ELECT|BAR\.|ACQBULK|FADD2|HMNMX2", l)]
            with open(fn, "w") as fh:
                fh.write(f"// {names[mangled]}\n// cuobjdump --dump-sass of libdml_b200.so (sm_100a), reduced to the tensor-core / TMEM / TMA / mbarrier / "
                         f"barrier instructions ({len(keep)} of {len(lines)} instructions); counts in {args.tag}_sass_summary.csv\n")
                fh.write("\n".join(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in keep) + "\n")
    with open(os.path.join(args.out_dir, f"{args.tag}_sass_summary.csv"), "w") as fh:
        fh.write("# per kernel of libdml_b200.so: instruction count and counts of the mnemonics that identify the tcgen05 / TMEM / TMA path\n")
        fh.write("kernel,template_args,instructions," + ",".join(KEYS) + "\n")
        for base, targs, n, c in sorted(rows):
            fh.write(f"{base},{targs},{n}," + ",".join(str(c[k]) for k in KEYS) + "\n")
    print(f"{len(rows)} kernels; summary in {args.out_dir}/{args.tag}_sass_summary.csv")


if __name__ == "__main__":
    sys.exit(main())
