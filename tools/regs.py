"""Developer probe: registers / spills per kernel from an nvcc -Xptxas -v log (default csrc/build.log)."""
import re
import subprocess
import sys

path = sys.argv[1] if len(sys.argv) > 1 else "raytracing_test_b200/csrc/build.log"
name, sp = None, ""
for l in open(path):
    m = re.search(r"Compiling entry function '([^']+)'", l)
    if m:
        name = m.group(1)
    if "spill" in l and name:
        sp = l.strip()
    m = re.search(r"Used (\d+) registers(.*)", l)
    if m and name:
        dn = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        dn = re.sub(r"\(.*", "", dn)[:70]
        print(f"{dn:70s} regs {m.group(1):>3s} | {sp}")
        name = None
