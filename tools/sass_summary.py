"""profiles/rNN_sass_summary.txt and rNN_ptxas.txt from the built library (no GPU needed):
    python tools/sass_summary.py r02
cuobjdump -sass mnemonic counts per kernel that identify the data path, and the -Xptxas -v register / shared-memory table."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
so = os.path.join(ROOT, "llm-d_b200", "b200ep", "libb200ep.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
blocks = re.split(r"\n\s*Function : \S+\n", sass)[1:]
pats = [("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("LDG.256", r"\bLDG\.[A-Z0-9.]*\b256"), ("STG.256", r"\bSTG\.[A-Z0-9.]*\b256"),
        ("F2FP.E4M3", r"F2FP[A-Z.]*E4M3"), ("F32x2", r"\b(FFMA2|FMUL2|FADD2)\b"), ("MEMBAR.SYS", r"MEMBAR[A-Z.]*SYS"),
        ("tensor", r"\b(UTC\w*MMA|LDTM|STTM|HMMA|IMMA|QMMA)\b")]
rows = []
for name, body in zip(names, blocks):
    short = re.sub(r"\(.*", "", name).strip()
    n_instr = len(re.findall(r"^\s+/\*[0-9a-f]{4}\*/", body, flags=re.M))
    rows.append((short, n_instr, [len(re.findall(p, body)) for _, p in pats]))
rows.sort()
with open(os.path.join(ROOT, "profiles", f"{tag}_sass_summary.txt"), "w") as f:
    f.write("cuobjdump -sass of libb200ep.so (sm_100a): counts of the mnemonics that identify the data path, per kernel\n"
            "UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier, LDG/STG ...256 = 32-byte accesses, F2FP..E4M3 = fp32 -> e4m3, "
            "FFMA2/FMUL2 = packed fp32,\nMEMBAR..SYS = system fence; tensor-core mnemonics (UTC*MMA, LDTM, HMMA): none, this path has no contraction\n\n")
    for short, n, c in rows:
        f.write(f"{short:<72} instr {n:5d}  " + "  ".join(f"{k} {v:3d}" for (k, _), v in zip(pats, c)) + "\n")
# ptxas table from the build logs
out = []
for log in sorted(os.listdir(os.path.join(ROOT, "llm-d_b200", "csrc"))):
    if not log.endswith(".ptxas.log"):
        continue
    txt = open(os.path.join(ROOT, "llm-d_b200", "csrc", log)).read()
    for m in re.finditer(r"Function properties for (\S+)\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n"
                         r"ptxas info\s+: Used (\d+) registers(?:, used (\d+) barriers)?(?:, (\d+) bytes cumulative stack size)?(?:, (\d+) bytes smem)?", txt):
        out.append((m.group(1), int(m.group(5)), int(m.group(3)) + int(m.group(4)), int(m.group(8) or 0), log[:-10]))
dem = subprocess.run(["cu++filt"], input="\n".join(o[0] for o in out), capture_output=True, text=True).stdout.split("\n")
with open(os.path.join(ROOT, "profiles", f"{tag}_ptxas.txt"), "w") as f:
    f.write("nvcc -Xptxas -v (sm_100a), per kernel: registers / spill bytes / static shared memory\n\n")
    for (_, regs, spill, smem, src), name in sorted(zip(out, dem), key=lambda t: t[1]):
        short = name.split("(")[0].strip()
        f.write(f"{short:<72} {src + '.cu':<12} regs {regs:3d}  spill {spill:3d}  smem {smem:6d}\n")
print("wrote", tag)
