#!/usr/bin/env python3
"""SASS evidence for every kernel of libddlb_b200.so (no GPU needed): an index with the
mnemonics that identify a Blackwell-native fused kernel, plus gzip'ed full listings of the
flagship instantiations. Writes profiles/sass/."""
import gzip
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, "ddlb_b200", "_native", "libddlb_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCHMMA.2CTA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "UTCATOMSWS",
        "LDGMC", "MULTIMEM", "RED.E.ADD", "REDG", "LDG.E.128", "STG.E.128", "LDG.E.EF.128", "STG.E.EF.128",
        "MEMBAR.ALL.SYS", "MEMBAR.SC.SYS", "CCTL.IVALL", "HMMA", "NANOSLEEP"]
FLAGSHIP = ("tp_gemm_kernelILi2ELi256ELb1ELb0ELb0ELb0ELi1E", "tp_gemm_kernelILi2ELi256ELb0ELb1ELb1ELb0ELi1E",
            "tp_gemm_kernelILi2ELi256ELb1ELb0ELb0ELb0ELi2E")


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
        if m and cur:
            usage[cur] = (int(m.group(1)), int(m.group(2)))
    parts = re.split(r"\n\s*Function : ", txt)
    rows = []
    for part in parts[1:]:
        name, _, body = part.partition("\n")
        name = name.strip()
        lines = [re.sub(r"/\*[0-9a-fA-Fx]+\*/", "", ln).strip() for ln in body.splitlines()]
        instr = [ln for ln in lines if ln and not ln.startswith(".") and not ln.startswith("/*") and ";" in ln]
        counts = {k: sum(1 for ln in instr if re.search(r"(^|\s)" + re.escape(k) + r"(\.|\s|;)", ln)) for k in KEYS}
        demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
        rows.append((demangled, name, len(instr), counts, usage.get(name, ("?", "?"))))
        if any(f in name for f in FLAGSHIP):
            path = os.path.join(OUT, re.sub(r"[^A-Za-z0-9_]", "_", name)[:120] + ".sass.gz")
            with open(path, "wb") as raw, gzip.GzipFile(fileobj=raw, mode="wb", filename="", mtime=0) as fh:
                fh.write(("// " + demangled + "\n" + "\n".join(lines)).encode())   # mtime 0: reproducible
    show = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "LDGMC", "REDG", "LDG.E.EF.128", "STG.E.128",
            "MEMBAR.ALL.SYS", "CCTL.IVALL", "HMMA"]
    with open(os.path.join(OUT, "INDEX.md"), "w") as fh:
        fh.write("# SASS index of libddlb_b200.so (cuobjdump -sass, sm_100a)\n\n"
                 "`UTCHMMA` = tcgen05.mma (kind::f16/tf32), `UTMALDG/UTMASTG` = TMA tensor load/store, `LDTM` = tcgen05.ld, "
                 "`UTCBAR` = tcgen05.commit, `LDGMC` = multimem.ld_reduce, `REDG` = flag bumps, `LDG.E.EF.128` / "
                 "`STG.E.128` = register-path peer copies, `HMMA` = legacy mma.sync (must be 0). "
                 "Template arguments: <cta_group, BLOCK_N, B N-major, fp32, 3xTF32 split, [trace,] pairs>.\n\n")
        fh.write("| kernel | instructions | regs | " + " | ".join(show) + " |\n|---|---|---|" + "---|" * len(show) + "\n")
        for demangled, name, n, counts, (regs, _) in sorted(rows):
            short = demangled.replace("ddlb::", "").replace("(ddlb::TpGemmParams)", "").replace("(ddlb::TpGemmParams const*, int)", "")
            short = re.sub(r"\((int|bool)\)", "", short)
            fh.write(f"| `{short[:90]}` | {n} | {regs} | " + " | ".join(str(counts[k]) for k in show) + " |\n")
    print(f"{len(rows)} kernels -> {OUT}")


if __name__ == "__main__":
    sys.exit(main())
