"""Reproducible SASS evidence for the hot kernels (developer tool; runs in the build container).

    python tools/sass_summary.py [--lib waveforms_b200/_lib/libwfb200.so] [--out profiles/sass] [--all]

Runs `cuobjdump -res-usage` and `cuobjdump -sass` on the built library and writes, per kernel,
profiles/sass/<short name>.txt with

  * registers / stack (spill) / shared / constant usage,
  * an opcode histogram of the whole kernel (MUFU, FFMA2, FFMA, IMAD, LOP3, SHFL, LDS, LDGSTS, REDUX,
    STL/LDL, ...) and of its hottest loop (the longest backward branch span),
  * the SASS of that loop (so DESIGN.md's instruction counts can be checked without rebuilding),

plus profiles/sass/INDEX.md, one line per kernel.  _lib/ is git-ignored, so these text files are the
tracked record of what the compiler produced for the tree they are committed with.
"""
from __future__ import annotations

import argparse
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOT = ("encode_frames_kernel", "cpm_modulate_f32s8_kernel", "mf_bank_f32_kernel", "viterbi_soqpsk_l2_kernel",
       "viterbi_cpm", "cpm_detect", "cpm_modulate_bank_warp", "welch", "phase_tree", "timing_metric", "eye_traces",
       "awgn_kernel<float, true>")
WATCH = ["MUFU", "FFMA2", "FFMA", "FADD2", "FMUL2", "FADD", "FMUL", "FMNMX", "FSETP", "FSEL", "SEL", "IMAD", "IADD3", "LOP3",
         "SHF", "PRMT", "POPC", "IDP", "SHFL", "REDUX", "CREDUX", "VOTE", "LDS", "STS", "LDG", "STG", "LDGSTS", "LDGDEPBAR",
         "DEPBAR", "LDC", "LDCU", "ULDC", "UMOV", "MOV", "BAR", "WARPSYNC", "BRA", "LDL", "STL", "ATOM", "RED", "I2F", "F2I",
         "UTCMMA", "UTMALDG", "LDTM", "SYNCS"]


def run(cmd):
    return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, check=True).stdout


def demangle(names):
    out = run(["cu++filt", *names]).splitlines() if names else []
    return dict(zip(names, out))


def short(dem):
    s = re.sub(r"^void\s+", "", dem)
    s = re.sub(r"\((?:int|bool|unsigned int|long|unsigned long)\)", "", s)  # template value casts
    s = re.sub(r"\(.*$", "", s)
    s = s.replace("wf::", "")
    s = re.sub(r"[^A-Za-z0-9_<>,\-]+", "", s.replace(" ", ""))
    s = s.replace("<", "__").replace(">", "").replace(",", "_")
    return s[:120]


def opcode(line):
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_.]+)?)", line)
    return m.group(1) if m else None


def hist(lines):
    c = collections.Counter()
    for ln in lines:
        op = opcode(ln)
        if op:
            c[op.split(".")[0]] += 1
            if op.startswith("MUFU") or op.startswith("LDS") or op.startswith("STS") or op.startswith("SHFL"):
                c[op] += 1
    return c


def loops(lines):
    """All backward-branch spans (start, end) as line indices, innermost first by length."""
    addr = {}
    for i, ln in enumerate(lines):
        m = re.match(r"\s*/\*([0-9a-f]+)\*/", ln)
        if m:
            addr[int(m.group(1), 16)] = i
    out = set()
    for i, ln in enumerate(lines):
        m = re.search(r"BRA[A-Z.]*\s+(?:!?U?P\d+,\s*)?0x([0-9a-f]+)\s*;", ln)
        if opcode(ln) and opcode(ln).startswith("BRA") and m:
            tgt = int(m.group(1), 16)
            if tgt in addr and addr[tgt] < i:
                out.add((addr[tgt], i))
    return sorted(out, key=lambda s: s[1] - s[0])


WORK = ("MUFU", "FFMA2", "FFMA", "FADD", "FADD2", "FMUL", "FMUL2", "FSEL", "FSETP", "FMNMX", "SHFL", "LOP3", "IMAD", "SEL")


def hottest_loop(lines):
    """The SMALLEST loop that holds at least 60 % of the kernel's arithmetic instructions (the unrolled
    inner body the kernel spends its time in); the longest loop when no loop qualifies."""
    ls = loops(lines)
    if not ls:
        return None
    total = sum(v for k, v in hist(lines).items() if k in WORK)
    for sp in ls:
        if sum(v for k, v in hist(lines[sp[0]:sp[1] + 1]).items() if k in WORK) >= 0.6 * total:
            return sp
    return ls[-1]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default=os.path.join(ROOT, "waveforms_b200", "_lib", "libwfb200.so"))
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "sass"))
    ap.add_argument("--all", action="store_true", help="every kernel, not only the hot ones")
    ap.add_argument("--max-loop-lines", type=int, default=400)
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)

    res = {}
    cur = None
    for ln in run(["cuobjdump", "-res-usage", a.lib]).splitlines():
        m = re.match(r"\s*Function\s+(\S+):", ln)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in ln:
            res[cur] = ln.strip()
            cur = None
    sass = run(["cuobjdump", "-sass", a.lib])
    blocks = re.split(r"\n\s*Function : ", sass)[1:]
    names = [b.split("\n", 1)[0].strip() for b in blocks]
    dem = demangle(names)
    index = []
    for name, blk in zip(names, blocks):
        d = dem.get(name, name)
        if not a.all and not any(h in d for h in HOT):
            continue
        lines = [ln for ln in blk.splitlines() if re.match(r"\s*/\*[0-9a-f]{4}\*/", ln)]
        h = hist(lines)
        loop = hottest_loop(lines)
        path = os.path.join(a.out, short(d) + ".txt")
        with open(path, "w") as f:
            f.write(f"kernel   : {d}\nmangled  : {name}\nresources: {res.get(name, '?')}\n")
            f.write(f"instructions: {len(lines)}\n\nopcode histogram, whole kernel (watched opcodes):\n")
            f.write("  " + "  ".join(f"{k}={h[k]}" for k in WATCH if h.get(k)) + "\n")
            f.write("  MUFU by function: " + "  ".join(f"{k}={v}" for k, v in sorted(h.items()) if k.startswith("MUFU.")) + "\n")
            if loop:
                body = lines[loop[0]:loop[1] + 1]
                hl = hist(body)
                f.write("\nall loops (instructions): " + ", ".join(str(e - b + 1) for b, e in loops(lines)) + "\n")
                f.write(f"hot loop (smallest holding >= 60 % of the arithmetic): {len(body)} instructions (SASS lines {loop[0]}..{loop[1]})\n")
                f.write("  " + "  ".join(f"{k}={hl[k]}" for k in WATCH if hl.get(k)) + "\n\n")
                shown = body if len(body) <= a.max_loop_lines else body[: a.max_loop_lines // 2] + ["        ..."] + body[-a.max_loop_lines // 2:]
                f.write("\n".join(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln) for ln in shown) + "\n")
        spill = re.search(r"STACK:(\d+)", res.get(name, ""))
        index.append((short(d), res.get(name, "?"), len(lines), h.get("MUFU", 0), h.get("FFMA2", 0), h.get("LDGSTS", 0),
                      h.get("STL", 0) + h.get("LDL", 0), int(spill.group(1)) if spill else -1))
    with open(os.path.join(a.out, "INDEX.md"), "w") as f:
        f.write("# SASS summaries (tools/sass_summary.py; one file per kernel instance)\n\n")
        f.write("| kernel | resources | instr | MUFU | FFMA2 | LDGSTS | STL+LDL | stack B |\n|---|---|---|---|---|---|---|---|\n")
        for row in sorted(index):
            f.write("| " + " | ".join(str(x) for x in row) + " |\n")
    print(f"{len(index)} kernels -> {a.out}")


if __name__ == "__main__":
    main()
