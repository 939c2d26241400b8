"""SASS evidence per object file: counts of the Blackwell-native opcodes (tcgen05.mma -> UTC*MMA, tcgen05.ld/st -> LDTM/STTM,
TMA -> UTMALDG/UTMASTG/UTMAREDG/UBLKCP) and of the legacy tensor path (HMMA) in espnet_slurp_b200/build/*.o.

    python tools/sass_histogram.py > profiles/r2_sass_histogram.txt
"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PAT = re.compile(r"\b(UTC\w*MMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMA\w+[.\w]*|UBLKCP[.\w]*|HMMA[.\w]*|UTCBAR[.\w]*|SYNCS[.\w]*|REDUX[.\w]*|MUFU\.\w+)")

for o in sorted(glob.glob(os.path.join(ROOT, "espnet_slurp_b200", "build", "*.o"))):
    sass = subprocess.run(["cuobjdump", "-sass", o], capture_output=True, text=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    kernels = len(re.findall(r"^\s*Function : ", sass, re.M))
    cnt = collections.Counter(m.group(1) for m in PAT.finditer(sass))
    fam = collections.Counter()
    for k, v in cnt.items():
        fam[k.split(".")[0]] += v
    print(f"== {os.path.basename(o)}  arch {','.join(arch)}  kernels {kernels}")
    print("   " + "  ".join(f"{k}={v}" for k, v in sorted(fam.items())))
    detail = {k: v for k, v in cnt.items() if k.startswith(("UTC", "UTMA", "LDTM", "STTM"))}
    if detail:
        print("   detail: " + "  ".join(f"{k}={v}" for k, v in sorted(detail.items())))
