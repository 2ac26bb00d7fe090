"""Counts of the Blackwell-native SASS mnemonics per fused-GEMM kernel (cuobjdump -sass of libslider_b200.so).
  python tools/sass_summary.py > profiles/sass_summary.txt"""
import collections
import re
import subprocess
from pathlib import Path

so = str(Path(__file__).resolve().parent.parent / "sliders_conceptmod_b200" / "libslider_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
COLS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "UTCBAR", "UTCATOMSWS", "SYNCS", "HMMA"]
cnt, cur = collections.defaultdict(collections.Counter), None
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = m.group(1)
        continue
    if cur is None:
        continue
    m = re.search(r"\s([A-Z][A-Z0-9]+(?:\.\w+)*)\s", l)
    if not m:
        continue
    op = m.group(1)
    base = op.split(".")[0]
    if base == "UTCHMMA":
        cnt[cur]["UTCHMMA.2CTA" if ".2CTA" in op else "UTCHMMA"] += 1
        if ".2CTA" in op:
            cnt[cur]["UTCHMMA"] += 1
    elif base in COLS:
        cnt[cur][base] += 1
names = sorted(cnt)
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
print("# cuobjdump -sass sliders_conceptmod_b200/libslider_b200.so: occurrences of the Blackwell-native mnemonics per kernel")
print("# (tcgen05.mma -> UTCHMMA[.2CTA], tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG/UTMAREDG/UTMAPF, tcgen05.commit -> UTCBAR,")
print("#  tcgen05.alloc -> UTCATOMSWS, mbarrier -> SYNCS; HMMA would be the legacy mma.sync path: none in the GEMM)")
print("kernel | " + " ".join(COLS))
rows = []
for n, d in zip(names, dem):
    if "lora_gemm_kernel" not in d:
        continue
    rows.append((re.sub(r"\(.*", "", d).replace("void ", ""), " ".join(str(cnt[n][c]) for c in COLS)))
for r in sorted(rows):
    print(f"{r[0]} | {r[1]}")
