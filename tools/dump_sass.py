#!/usr/bin/env python3
"""profiles/sass/<kernel>.sass: `cuobjdump -sass` of the hot kernels of the built library, one file per kernel, with a
header line counting its tcgen05 / TMA / TMEM opcodes (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA
load / store, UTCBAR = tcgen05.commit).  Run after `make -C deepsdf_b200/csrc` (no GPU needed):  python tools/dump_sass.py"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "deepsdf_b200", "libdsdf_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
WANT = {"chain_forward_train": "chain_forward_kernelILb1", "chain_forward_eval": "chain_forward_kernelILb0",
        "chain_dgrad": "chain_dgrad_kernel", "wgrad": "_ZN4dsdf12wgrad_kernel", "input_grad": "input_grad_kernel",
        "small_forward_128": "small_forward_kernelILi128", "chain_exact": "chain_exact_kernel"}
SPECIAL = ("UTC", "UTMA", "LDTM", "SYNCS", "UBLKCP")


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    os.makedirs(OUT, exist_ok=True)
    for name, key in WANT.items():
        body = [p for p in parts if key in p.split("\n", 1)[0]][0]
        lines = body.split("\n")
        ops = collections.Counter()
        for ln in lines:
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if m:
                op = m.group(1)
                ops[".".join(op.split(".")[:3]) if op.startswith(SPECIAL) else op.split(".")[0]] += 1
        with open(os.path.join(OUT, name + ".sass"), "w") as f:
            f.write(f"// cuobjdump -sass deepsdf_b200/libdsdf_b200.so, function {lines[0].strip()}\n")
            f.write("// tcgen05 / TMA / TMEM opcodes: " + ", ".join(f"{k} x{v}" for k, v in sorted(ops.items()) if k.startswith(SPECIAL)) + "\n")
            f.write(f"// {sum(ops.values())} instructions\n")
            f.write("\n".join(lines[1:]))
        print(name, sum(ops.values()), "instructions")


if __name__ == "__main__":
    main()
