"""Summarise an .ncu-rep (read on the CPU box): per kernel time, occupancy, lanes/instruction, issue rate,
DRAM bytes, cache hit rates and the top warp-stall reasons.  usage: python tools/ncu_summary.py rep.ncu-rep"""
import csv, subprocess, sys, io

def main(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    keys = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
            "sm__inst_executed.sum", "sm__inst_executed.sum.per_cycle_active", "sm__inst_executed.sum.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__icc_request_hit_rate.pct"]
    for r in data:
        print("----", r[idx["Kernel Name"]][:60])
        for k in keys:
            if k in idx:
                print("    %-60s %-14s %s" % (k, units[idx[k]], r[idx[k]]))
        st = [(float(r[i].replace(",", "")), h) for h, i in idx.items()
              if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and "not_issued" not in h]
        print("    stalls per issue: " + ", ".join("%s %.2f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v)
                                                  for v, h in sorted(st, reverse=True)[:7]))

if __name__ == "__main__":
    main(sys.argv[1])
