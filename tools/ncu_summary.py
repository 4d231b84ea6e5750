"""Print a compact per-kernel summary of an .ncu-rep (raw page): duration, instructions, IPC, occupancy, DRAM bytes,
top stall reasons. Usage: python tools/ncu_summary.py report.ncu-rep [out.md]"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def g(r, k, default=""):
        return r[idx[k]] if k in idx else default

    out = []
    stall_keys = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    for r in rows[2:]:
        name = g(r, "Kernel Name").split("(")[0].replace("void ", "")
        stalls = sorted(((float(r[idx[k]]), k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
                         for k in stall_keys if r[idx[k]] not in ("", "n/a")), reverse=True)[:4]
        out.append(
            f"{name}: {g(r,'gpu__time_duration.sum')} {units[idx['gpu__time_duration.sum']]}, grid {g(r,'launch__grid_size')} x {g(r,'launch__block_size')}, "
            f"regs {g(r,'launch__registers_per_thread')}, inst {float(g(r,'smsp__inst_executed.sum','0'))/1e6:.1f}M, "
            f"issue-active {g(r,'smsp__issue_active.avg.pct_of_peak_sustained_active')}%, warps-active {g(r,'sm__warps_active.avg.pct_of_peak_sustained_active')}%, "
            f"thr/inst {g(r,'smsp__thread_inst_executed_per_inst_executed.ratio')}, "
            f"dram rd {g(r,'dram__bytes_read.sum')} {units[idx['dram__bytes_read.sum']]} wr {g(r,'dram__bytes_write.sum')} {units[idx['dram__bytes_write.sum']]}, "
            f"dram-thr {g(r,'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')}%, "
            f"l2-thr {g(r,'lts__throughput.avg.pct_of_peak_sustained_elapsed')}%, "
            f"sm-cycles max {g(r,'sm__cycles_elapsed.max')} active-avg {g(r,'smsp__cycles_active.avg')}, "
            f"stalls {', '.join(f'{n}={v:.2f}' for v, n in stalls)}")
    text = "\n".join(out)
    print(text)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text + "\n")


if __name__ == "__main__":
    main()
