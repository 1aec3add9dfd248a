"""DEVELOPMENT TOOL ONLY: compile the libwfb200 kernel sources for HOST threads.

    python tools/hostemu/build_emu.py  ->  /tmp/wfemu/libwfb200_emu.so

Used for pre-flight debugging of kernel logic in a container without a GPU (see
tools/hostemu/README.md).  The product never loads this library.
"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "waveforms_b200", "csrc")
OUT = os.environ.get("WFEMU_DIR", "/tmp/wfemu")

LAUNCH = re.compile(r"^(\s*)(else if \(.*?\) |if \(.*?\) |else )?([A-Za-z_][\w:<>, ]*?)<<<(.*?)>>>\((.*)\);(\s*\\?)$")


def split_top(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "([{<" and not (ch == "<" and False):
            depth += ch in "([{"
        if ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    out.append(cur.strip())
    return out


def convert(text):
    lines = text.split("\n")
    res = []
    i = 0
    while i < len(lines):
        ln = lines[i]
        # join a launch that spans two lines
        if "<<<" in ln and ">>>" in ln and not ln.rstrip().rstrip("\\").rstrip().endswith(";"):
            ln = ln.rstrip() + " " + lines[i + 1].strip()
            if not ln.rstrip().rstrip("\\").rstrip().endswith(";"):
                ln = ln.rstrip() + " " + lines[i + 2].strip()
                i += 1
            i += 1
        m = LAUNCH.match(ln)
        if m:
            ind, pre, kern, cfg, args, tail = m.groups()
            cfgs = split_top(cfg)
            while len(cfgs) < 4:
                cfgs.append("0")
            ln = f"{ind}{pre or ''}::emu::launch(dim3({cfgs[0]}), dim3({cfgs[1]}), (size_t)({cfgs[2]}), [&] {{ {kern.strip()}({args}); }});{tail}"
        elif "<<<" in ln:
            raise SystemExit(f"unconverted launch: {ln}")
        ln = ln.replace("extern __shared__ __align__(16) unsigned char smem_raw[];",
                        "unsigned char* smem_raw = ::emu::dyn_smem();")
        res.append(ln)
        i += 1
    return "\n".join(res)


def main():
    os.makedirs(OUT, exist_ok=True)
    srcs = []
    for name in sorted(os.listdir(CSRC)):
        path = os.path.join(CSRC, name)
        text = open(path).read()
        if name.endswith(".cu"):
            text = "namespace emu { unsigned char* dyn_smem(); }\n" + convert(text)
            dst = os.path.join(OUT, name[:-3] + ".cpp")
            srcs.append(dst)
        else:
            dst = os.path.join(OUT, name)
        open(dst, "w").write(text)
    inc_dir = os.path.join(OUT, "include")
    os.makedirs(inc_dir, exist_ok=True)
    # sources include "../../include/wfb200.h" relative to csrc: mirror that layout
    os.makedirs(os.path.join(OUT, "..", "..", "include"), exist_ok=True)
    lib = os.path.join(OUT, "libwfb200_emu.so")
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fPIC", "-shared", "-pthread", "-w",
           "-D__CUDACC__=1", "-I", HERE, "-I", OUT,
           "-include", os.path.join(HERE, "cuda_runtime.h"),
           os.path.join(HERE, "emu.cpp"), *srcs, "-o", lib]
    hdr = os.path.join(ROOT, "include", "wfb200.h")
    for s in srcs + [os.path.join(OUT, "common.cuh")]:
        t = open(s).read().replace('"../../include/wfb200.h"', f'"{hdr}"')
        open(s, "w").write(t)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout[-6000:])
        raise SystemExit("hostemu build failed")
    print(lib)


if __name__ == "__main__":
    main()
