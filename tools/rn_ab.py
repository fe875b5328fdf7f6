"""A/B of two builds of the native library on the ResNet-18 step (per-layer conv times):
    python tools/rn_ab.py nanofed_b200/_native/libnanofed_b200_prev.so nanofed_b200/_native/libnanofed_b200.so
Each library runs `tools/rn_prof.py` in its own process (NANOFED_B200_LIB), twice, alternating."""
import os, re, subprocess, sys
from collections import defaultdict
libs = sys.argv[1:3]
res = {l: defaultdict(list) for l in libs}
steps = {l: [] for l in libs}
for rep in range(2):
    for l in libs:
        env = dict(os.environ, NANOFED_B200_LIB=os.path.abspath(l))
        out = subprocess.run([sys.executable, "tools/rn_prof.py", "--reps", "2"], env=env,
                             capture_output=True, text=True).stdout
        for line in out.splitlines():
            m = re.match(r"\s*([\d.]+) us\s+[\d.]+%\s+x(\d+)\s+(conv_\w+ \S+ @\S+ taps\d+)", line)
            if m:
                res[l][m.group(3)].append(float(m.group(1)) / int(m.group(2)))
            m = re.match(r"local step.*?: ([\d.]+) ms eager", line)
            if m:
                steps[l].append(float(m.group(1)))
print("step ms:", {os.path.basename(l): steps[l] for l in libs})
for name in sorted(res[libs[0]]):
    a, b = min(res[libs[0]][name]), min(res[libs[1]].get(name, [float("nan")]))
    print(f"{name:42s} {a:8.1f} -> {b:8.1f} us  ({b / a:.2f}x)")
