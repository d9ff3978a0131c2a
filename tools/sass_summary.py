"""Per-kernel SASS opcode summary of libdpvo_b200.so (cuobjdump -sass): counts of the mnemonics that tell which
hardware path a kernel uses — tcgen05 (UTC*MMA, LDTM/STTM), TMA (UTMALDG/UTMASTG/UBLKCP), legacy tensor (HMMA),
async copies (LDGSTS), wide loads (LDG.E.128/256), reductions/atomics (RED/ATOM/ATOMS).
    python tools/sass_summary.py > profiles/r02_sass_opcodes.json"""
import collections
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "deep-visual-odometry-slam_b200", "libdpvo_b200.so")
PAT = [("UTCHMMA", r"\bUTCHMMA"), ("UTCQMMA", r"\bUTCQMMA"), ("UTCMMA_other", r"\bUTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"),
       ("STTM", r"\bSTTM"), ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"), ("UBLKCP", r"\bUBLKCP"),
       ("HMMA", r"\bHMMA"), ("LDGSTS", r"\bLDGSTS"), ("LDG.256", r"\bLDG\.E[.\w]*\.256"), ("LDG.128", r"\bLDG\.E[.\w]*\.128"),
       ("REDG", r"\bREDG?\."), ("REDG.F32x4", r"\bREDG\.E\.ADD\.F32x4"), ("REDG.F64", r"\bREDG\.E\.ADD\.F64"),
       ("ATOM", r"\bATOM\."), ("ATOMS", r"\bATOMS\."), ("ATOMG", r"\bATOMG\."),
       ("LDL", r"\bLDL"), ("STL", r"\bSTL"), ("BAR", r"\bBAR\."), ("SHFL", r"\bSHFL\."), ("DFMA", r"\bDFMA"), ("FFMA", r"\bFFMA")]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    out = collections.OrderedDict()
    name = None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            name = re.sub(r"\(.*", "", name)
            out[name] = collections.Counter()
            out[name]["instructions"] = 0
            continue
        if name and re.search(r"/\*[0-9a-f]{4,}\*/", line):
            out[name]["instructions"] += 1
            for key, pat in PAT:
                if re.search(pat, line):
                    out[name][key] += 1
    res = {k: {a: b for a, b in v.items() if b} for k, v in out.items()}
    json.dump(res, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
