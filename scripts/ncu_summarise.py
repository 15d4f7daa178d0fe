"""Reads gpurun_out/ncu_*.ncu-rep (no GPU needed) and writes profiles/ncu_<kernel>.txt: duration, DRAM
bytes and throughput, tensor-pipe and issue utilisation, registers, occupancy, top stall reasons."""
import csv
import glob
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_l1tex2xbar_write_bytes.sum"]


def main():
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    for rep in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "ncu_*.ncu-rep")) + sys.argv[1:]):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        if len(rows) < 3:
            print("empty report", rep)
            continue
        hdr, units, val = rows[0], rows[1], rows[2]
        name = val[hdr.index("Kernel Name")] if "Kernel Name" in hdr else os.path.basename(rep)
        out = [f"ncu --set full --clock-control none, one launch of {name}", f"report: gpurun_out/{os.path.basename(rep)} (scratch, not committed)", ""]
        for i, h in enumerate(hdr):
            if any(h == k or h.startswith(k) for k in KEYS):
                out.append(f"{h:75s} {val[i]:>16s} {units[i]}")
        stalls = [(float(val[i] or 0), h) for i, h in enumerate(hdr)
                  if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") and val[i]]
        if not stalls:
            stalls = [(float(val[i] or 0), h) for i, h in enumerate(hdr) if "issue_stalled" in h and h.endswith(".pct") and val[i]]
        out.append("")
        out.append("top stall reasons:")
        for v, h in sorted(stalls, reverse=True)[:6]:
            out.append(f"  {v:10.3f}  {h}")
        dst = os.path.join(ROOT, "profiles", os.path.basename(rep).replace(".ncu-rep", ".txt"))
        open(dst, "w").write("\n".join(out) + "\n")
        print("wrote", dst)


if __name__ == "__main__":
    main()
