"""Per-source-line stall samples of one kernel from an .ncu-rep (run where ncu/nvdisasm exist).

    python scripts/ncu_by_line.py gpurun_out/attn_ncu.ncu-rep attn_bwd_dkdv_kernel attention [top]

`ncu --page source --csv` lists SASS instructions with sampling counters but no line numbers; the
line table comes from `nvdisasm -g` of the cubin extracted from the in-tree .so (built with
-lineinfo).  The two are joined on the instruction offset inside the function.
"""
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "distributed-training-and-deepspeed_b200", "lib", "libdtds_b200.so")


def line_table(cubin_key: str, kernel: str):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", cubin_key, SO], cwd=tmp, check=True, capture_output=True)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    table, cur, inside = {}, None, False
    for ln in dis.splitlines():
        if ln.startswith(".text."):
            inside = kernel in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            table[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return table


def main():
    rep, kernel, key = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kernel],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    idx = {n: i for i, n in enumerate(hdr)}
    data = []
    for r in rows[h + 1:]:
        if len(r) != len(hdr):
            break   # next kernel instance
        data.append(r)
    base = int(data[0][0], 16)
    table = line_table(key, kernel)
    stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]

    def f(x):
        try:
            return float(x)
        except ValueError:
            return 0.0
    per_line, total = {}, 0.0
    for r in data:
        off = int(r[0], 16) - base
        loc, _ = table.get(off, (None, ""))
        s = f(r[idx["# Samples"]])
        total += s
        d = per_line.setdefault(loc, {"samples": 0.0, "inst": 0.0, **{k: 0.0 for k in stalls}})
        d["samples"] += s
        d["inst"] += f(r[idx["Instructions Executed"]])
        for k in stalls:
            d[k] += f(r[idx[k]])
    print(f"{kernel}: {len(data)} SASS instructions, {int(total)} samples")
    src_cache = {}
    for loc, d in sorted(per_line.items(), key=lambda kv: -kv[1]["samples"])[:top]:
        best = sorted(((k[6:], v) for k, v in d.items() if k.startswith("stall_")), key=lambda kv: -kv[1])[:3]
        text = ""
        if loc:
            path = os.path.join(ROOT, "distributed-training-and-deepspeed_b200", "csrc", loc[0])
            if path not in src_cache and os.path.exists(path):
                src_cache[path] = open(path).read().splitlines()
            if path in src_cache and loc[1] <= len(src_cache[path]):
                text = src_cache[path][loc[1] - 1].strip()[:80]
        print(f"{100 * d['samples'] / max(total, 1):5.1f}%  {loc[0] if loc else '?'}:{loc[1] if loc else 0:<5d} "
              f"inst {int(d['inst']):>8d}  {[(k, int(v)) for k, v in best if v]}  | {text}")


if __name__ == "__main__":
    main()
