"""Per-kernel hot spots from an .ncu-rep source page: shared-memory wavefront excess, global
sector excess and stall samples per SASS instruction.

    python tools/ncu_hotspots.py gpurun_out/x.ncu-rep [kernel-regex] [top_n]
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
regex = sys.argv[2] if len(sys.argv) > 2 else ".*"
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 8
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name",
                      f"regex:{regex}"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(raw)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": {}}
        blocks.append(cur)
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
    elif cur is not None and cur["hdr"] and len(row) == len(cur["hdr"]):
        cur["rows"][row[0]] = row          # ncu prints every line twice: keep one per address


def num(row, ix, key):
    try:
        return int(float(row[ix[key]]))
    except (KeyError, ValueError):
        return 0


for b in blocks:
    ix = {h: i for i, h in enumerate(b["hdr"])}
    rows = list(b["rows"].values())
    inst = sum(num(r, ix, "Instructions Executed") for r in rows)
    wf = sum(num(r, ix, "L1 Wavefronts Shared") for r in rows)
    ex = sum(num(r, ix, "L1 Wavefronts Shared Excessive") for r in rows)
    gs = sum(num(r, ix, "L2 Theoretical Sectors Global") for r in rows)
    gx = sum(num(r, ix, "L2 Theoretical Sectors Global Excessive") for r in rows)
    print(f"## {b['name'][:100]}\n   warp instructions {inst}, smem wavefronts {wf} (excess {ex}), "
          f"global sectors {gs} (excess {gx})")
    for title, key in (("smem excess", "L1 Wavefronts Shared Excessive"),
                       ("global sector excess", "L2 Theoretical Sectors Global Excessive"),
                       ("stall samples", "# Samples")):
        best = sorted(rows, key=lambda r: -num(r, ix, key))[:top_n]
        best = [r for r in best if num(r, ix, key) > 0]
        if not best:
            continue
        print(f"   top by {title}:")
        for r in best:
            print(f"     {num(r, ix, key):>10}  x{num(r, ix, 'Instructions Executed'):<8} {r[ix['Source']].strip()[:80]}")
