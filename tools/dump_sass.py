"""Write per-kernel SASS evidence into profiles/sass/: the tcgen05 / TMA / TMEM / multimem
mnemonic counts of every kernel in the built library plus full listings of the hot ones."""
import re
import subprocess
import sys
from collections import Counter
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "nanofed_b200" / "_native" / "libnanofed_b200.so"
OUT = ROOT / "profiles" / "sass"
OUT.mkdir(parents=True, exist_ok=True)
sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
kernels, cur = {}, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kernels[cur] = []
    elif cur:
        kernels[cur].append(line)
KEYS = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "LDG.E.128",
        "STG.E.128", "ATOM", "RED", "LDGMC", "STGMC", "REDGMC", "MULTIMEM", "LD.E.STRONG.SYS", "ST.E.STRONG.SYS"]
rows = []
for name, lines in kernels.items():
    ops = Counter()
    for l in lines:
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m:
            for k in KEYS:
                if m.group(1).startswith(k) or k in m.group(1):
                    ops[k] += 1
    short = re.sub(r"\(.*", "", name)
    rows.append((short, len(lines), dict(ops)))
    if any(s in name for s in ("tc_conv", "tc_fc_dgrad", "tc_gemm_kernel<false, 128, false, 128",
                               "tc_gemm_kernel<true, 128, true, 128, 64, 128, 1, 2, 2, 1>",
                               "fedavg_allreduce_kernel<float>", "fedavg_reduce_kernel<float>", "fc_head",
                               "rn_conv_gemm_kernel<false, 128, 2>", "rn_conv_gemm_kernel<true, 64, 2>",
                               "rn_conv_wgrad_kernel<256, 128>", "tc_gemm_kernel<true, 128, true, 128, 64, 128, 1, 2, 2, 2>")):
        fn = re.sub(r"[^A-Za-z0-9_]+", "_", short)[:80]
        (OUT / f"{fn}.sass").write_text("\n".join(lines))
with open(OUT / "SUMMARY.md", "w") as f:
    f.write("# SASS mnemonic counts per kernel (cuobjdump -sass of libnanofed_b200.so)\n\n")
    f.write("`UTCHMMA` = tcgen05.mma, `UTMALDG`/`UTMASTG` = TMA load/store, `LDTM` = tcgen05.ld, "
            "`LDGMC`/`STGMC` = NVLS multimem.ld_reduce / multimem.st.*, `SYNCS` = mbarrier ops.\n\n| kernel | SASS lines | mnemonics |\n|---|---:|---|\n")
    for short, n, ops in sorted(rows):
        f.write(f"| `{short[:90]}` | {n} | {', '.join(f'{k}×{v}' for k, v in sorted(ops.items())) or '-'} |\n")
print(f"wrote {len(rows)} kernels to {OUT}")
