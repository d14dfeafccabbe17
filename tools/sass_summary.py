#!/usr/bin/env python
"""Evidence that the hot path is Blackwell-native, from the built objects (no GPU needed):
per object file the counts of the SASS mnemonics UTC*MMA (= tcgen05.mma; H f16/bf16, I int8, Q f8f6f4 / block
scaled), LDTM (tcgen05.ld), UTCCP (tcgen05.cp), UTMALDG (TMA load), .2CTA (cta_group::2) and the legacy
mma.sync forms, then registers / spills / shared memory per kernel from the ptxas -v build logs.
  python tools/sass_summary.py > profiles/sass_r2.txt      (after `python -m sgl_cpu_tests_b200.build`)"""
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "sgl_cpu_tests_b200", "build")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")
PATS = [("UTCHMMA", r"UTCHMMA"), ("UTCIMMA", r"UTCIMMA"), ("UTCQMMA", r"UTCQMMA"), (".2CTA", r"\.2CTA"), ("LDTM", r"LDTM"),
        ("UTCCP", r"UTCCP"), ("UTMALDG", r"UTMALDG"), ("UBLKCP", r"UBLKCP"), ("mma.sync", r"[IH]MMA\.16")]


def main():
    print("SASS mnemonic counts per object (cuobjdump -sass), sm_100a")
    print(f"{'object':22s}" + "".join(f"{n:>10s}" for n, _ in PATS))
    for o in sorted(glob.glob(os.path.join(BUILD, "*.o"))):
        sass = subprocess.run([os.path.join(CUDA, "bin", "cuobjdump"), "-sass", o], capture_output=True, text=True).stdout
        print(f"{os.path.basename(o)[:-2]:22s}" + "".join(f"{len(re.findall(p, sass)):10d}" for _, p in PATS))
    print("\nptxas -v per kernel: registers, spill bytes, static shared memory")
    rows = set()
    for log in sorted(glob.glob(os.path.join(BUILD, "*.o.log"))):
        text = open(log).read()
        for m in re.finditer(r"Compiling entry function '([^']+)' for 'sm_100a'\n(?:.*\n)*?.*?(\d+) bytes spill stores, (\d+) bytes spill loads\n"
                             r".*Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes smem)?", text):
            name = subprocess.run([os.path.join(CUDA, "bin", "cu++filt"), m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name[:name.index(">(") + 1] if ">(" in name else name.split("(")[0]
            name = name.replace("(int)", "").replace("(bool)", "").replace("sglb::", "").replace("<unnamed>::", "").replace("void ", "")
            rows.add(f"{name[:72]:74s} regs {int(m.group(4)):3d}  spill st/ld {m.group(2)}/{m.group(3)} B  smem {m.group(5) or 0} B")
    print("\n".join(sorted(rows)))


if __name__ == "__main__":
    main()
