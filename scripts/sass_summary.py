"""Regenerates profiles/sass_summary.txt: per-kernel counts of the SASS mnemonics that prove the
Blackwell paths are in the shipped library (cuobjdump -sass on lib/libdtds_b200.so)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "distributed-training-and-deepspeed_b200", "lib", "libdtds_b200.so")
PAT = re.compile(r"\b(UTCHMMA\S*|UTCBAR\S*|LDTM\S*|STTM\S*|UTMALDG\S*|UTMASTG\S*|SYNCS\S*|LDGMC\S*|STGMC\S*|REDGMC\S*|"
                 r"\S+\.STRONG\.SYS\S*|MEMBAR\S*|ACQBULK\S*|REDG\S*STRONG\.GPU\S*|LDG\S*STRONG\.GPU\S*)")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    per, cur = collections.OrderedDict(), None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?\S+\s+)?(\S+)", ln)
        if m and PAT.match(m.group(1)):
            per[cur][m.group(1).rstrip(";")] += 1
    out = ["SASS mnemonic counts per kernel of lib/libdtds_b200.so (cuobjdump -sass, sm_100a).",
           "UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG = TMA load, SYNCS = mbarrier, "
           "LDGMC/..MC = multimem (NVLS).", ""]
    for k, c in per.items():
        if not c:
            continue
        out.append(k)
        out.append("    " + ", ".join(f"{m} x{n}" for m, n in sorted(c.items())))
    open(os.path.join(ROOT, "profiles", "sass_summary.txt"), "w").write("\n".join(out) + "\n")
    print(len(per), "kernels;", sum(1 for c in per.values() if c.get("UTCHMMA") or any(k.startswith("UTCHMMA") for k in c)),
          "with tcgen05.mma")


if __name__ == "__main__":
    main()
