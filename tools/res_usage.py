"""Registers / spill stack / static shared memory of every kernel in libslider_b200.so (cuobjdump -res-usage), plus
the mma.sync (HMMA) and ldmatrix (LDSM) counts of the kernels that use the legacy tensor-core path on purpose (the
LoRA-gradient reductions and the short-key attention core: tiles far below a tcgen05 128-row atom).
  python tools/res_usage.py > profiles/r02_res_usage.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

so = str(Path(__file__).resolve().parent.parent / "sliders_conceptmod_b200" / "libslider_b200.so")
out = subprocess.run(["cuobjdump", "-res-usage", so], capture_output=True, text=True).stdout.splitlines()
names = [l.split("Function ")[1].rstrip(":") for l in out if "Function " in l]
usage = [l for l in out if l.strip().startswith("REG:")]
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
rows = []
for n, u in zip(dem, usage):
    f = dict(kv.split(":") for kv in u.split() if ":" in kv)
    rows.append((re.sub(r"\(.*", "", n).replace("void ", ""), int(f["REG"]), int(f["STACK"]), int(f["SHARED"]), int(f["LOCAL"])))
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
hm, ld, cur = collections.Counter(), collections.Counter(), None
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = m.group(1)
    elif cur:
        if " HMMA." in l:
            hm[cur] += 1
        elif " LDSM." in l:
            ld[cur] += 1
mang = dict(zip(dem, names))
print(f"# cuobjdump -res-usage / -sass of libslider_b200.so ({len(rows)} kernels); spill = STACK bytes; dynamic shared memory not shown")
print("kernel | registers | spill stack B | static smem B | HMMA | LDSM")
spills = 0
for n, reg, stack, sh, loc in sorted(rows):
    full = next((d for d in dem if d.replace("void ", "").startswith(n)), None)
    k = mang.get(full, "")
    print(f"{n} | {reg} | {stack} | {sh} | {hm.get(k, 0)} | {ld.get(k, 0)}")
    spills += stack > 0
print(f"# kernels with a spill stack: {spills}")
