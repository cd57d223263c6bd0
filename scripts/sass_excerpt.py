"""SASS evidence for the shipped kernels: `cuobjdump -sass` of libtstwo_cuda.so, per kernel the instruction mix (static
counts), the Blackwell-specific mnemonics it contains (TMA: UTMALDG/UTMASTG/UBLKCP, tensor cores: UTCIMMA/LDTM/STTM,
mbarrier: SYNCS) and an excerpt of its hottest basic block (the longest run of butterfly / round instructions).
usage: python scripts/sass_excerpt.py [kernel-name-regex ...] > profiles/rNN_sass.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tstwo_b200", "libtstwo_cuda.so")
DEFAULT = [r"fft_tma_kernel", r"merkle_leaf_tma_kernel", r"merkle_layer_kernel", r"merkle_leaf_absorb_kernel", r"quotients_mma_kernel",
           r"widefib_mma_kernel", r"fold_line", r"fold_circle", r"eval_at_point", r"bit_reverse"]
SPECIAL = re.compile(r"^(UTMA|UBLKCP|UTC|LDTM|STTM|SYNCS|IMMA|HMMA|LDGSTS|REDUX|ELECT)")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    pats = [re.compile(p) for p in (sys.argv[1:] or DEFAULT)]
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    funcs, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur:
            funcs[cur].append((m.group(1), m.group(2).strip()))
    names = demangle(list(funcs))
    print(f"# cuobjdump -sass {os.path.relpath(SO, ROOT)} (sm_100a); static instruction counts per kernel, Blackwell-specific")
    print("# mnemonics, and the first 28 instructions of the longest straight-line block of the kernel (its hot loop body).")
    for mangled, ins in funcs.items():
        nice = names.get(mangled, mangled)
        if not any(p.search(nice) for p in pats):
            continue
        short = re.sub(r"\(.*", "", nice.replace("(anonymous namespace)::", ""))
        mix = collections.Counter()
        for _, text in ins:
            ops = [o for o in text.split() if not o.startswith("@")]
            op = ops[0] if ops else "?"
            key = ".".join(op.split(".")[:2]) if op.startswith(("IMAD", "LEA", "LDS", "STS", "LDG", "STG", "VIADDMNMX", "IMMA", "UTC", "UTMA", "UBLKCP", "LDTM", "STTM", "SYNCS")) else op.split(".")[0]
            mix[key] += 1
        print("=" * 110)
        print(f"{short}   [{len(ins)} instructions]")
        print("  mix: " + ", ".join(f"{k} {v}" for k, v in mix.most_common(14)))
        sp = {k: v for k, v in mix.items() if SPECIAL.match(k)}
        if sp:
            print("  sm_100a-specific: " + ", ".join(f"{k} {v}" for k, v in sorted(sp.items())))
        # longest run without control flow
        best, start, run = (0, 0), 0, 0
        for i, (_, text) in enumerate(ins):
            if re.search(r"\b(BRA|EXIT|BAR|BSYNC|BSSY|CALL|RET|WARPSYNC|SYNCS|BREAK)\b", text):
                if run > best[0]:
                    best = (run, start)
                start, run = i + 1, 0
            else:
                run += 1
        if run > best[0]:
            best = (run, start)
        print(f"  longest straight-line block: {best[0]} instructions at /*{ins[best[1]][0]}*/; first 28:")
        for addr, text in ins[best[1]:best[1] + 28]:
            print(f"      /*{addr}*/  {text}")


if __name__ == "__main__":
    main()
