"""profiles/r02_sass_opcodes.txt: per-kernel SASS opcode histogram of the shipped objects (cuobjdump -sass on
ppo-libtorch_b200/csrc/build/*.o): the Blackwell-native mnemonics (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG
= cp.async.bulk.tensor, UBLKCP) and the CUDA-core mix the instruction-count work looks at (MUFU.*, F2FP, HFMA2, HMUL2 ...)."""
import collections
import glob
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "MUFU.TANH", "MUFU.EX2", "MUFU.RCP", "MUFU.LG2",
        "F2FP", "HFMA2", "HMUL2", "FFMA", "FMUL", "FADD", "DADD", "LDS", "STS", "LDG", "STG", "LDL", "STL", "SHFL", "BAR"]


def main():
    out = []
    for obj in sorted(glob.glob(os.path.join(ROOT, "ppo-libtorch_b200", "csrc", "build", "*.o"))):
        txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
        fn, hist = None, collections.OrderedDict()
        for line in txt.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
                hist[fn] = collections.Counter()
                continue
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)", line)
            if m and fn:
                op = m.group(1)
                hist[fn]["_total"] += 1
                for k in KEYS:
                    if op == k or op.startswith(k + "."):
                        hist[fn][k] += 1
        for fn, h in hist.items():
            if h["_total"] < 200:
                continue
            row = " ".join(f"{k}={h[k]}" for k in KEYS if h[k])
            out.append(f"{os.path.basename(obj):18s} {fn[:70]:70s} total={h['_total']:6d}  {row}")
    print("# cuobjdump -sass opcode counts per kernel (static), objects of", "this build")
    print("\n".join(out))


if __name__ == "__main__":
    sys.exit(main())
