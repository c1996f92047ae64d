"""Record the toolkit and a hash of the pop kernel's SASS (the build the shared-memory ordering of the fast pop
variant was verified on): python tools/sass_fingerprint.py > profiles/r02_pop_sass_fingerprint.txt"""
import hashlib, re, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
obj = ROOT / "flashgmm_b200/build/gmm_pop.o"
ver = subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout.strip().splitlines()[-2:]
sass = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
print("toolkit:", " | ".join(ver))
for m in re.finditer(r"Function : (\S+)\n(.*?)(?=\n\s*Function :|\Z)", sass, re.S):
    name, body = m.group(1), m.group(2)
    ops = [re.sub(r"/\*[0-9a-f]+\*/", "", l).strip() for l in body.splitlines() if "/*" in l and ";" in l]
    ops = [re.sub(r"\s+", " ", o) for o in ops]
    h = hashlib.sha256("\n".join(ops).encode()).hexdigest()[:16]
    n_sts = sum("STS" in o for o in ops); n_lds = sum("LDS" in o for o in ops); n_warpsync = sum("WARPSYNC" in o or "BAR" in o for o in ops)
    short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    print(f"{h}  insts={len(ops):5d} STS={n_sts:4d} LDS={n_lds:4d} WARPSYNC/BAR={n_warpsync:3d}  {short[:110]}")
