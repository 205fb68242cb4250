"""Key metrics of every kernel in an ncu report.  usage: python tools/ncu_summary.py report.ncu-rep"""
import csv
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__cluster_size",
        "launch__shared_mem_per_block_dynamic", "inst_executed", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "smsp__average_warp_latency_per_inst_issued.ratio", "sm__cycles_active.avg",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:100])
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print("  %-66s %s %s" % (w, r[i], units[i]))
    st = [(float(r[i].replace(",", "")), h) for i, h in enumerate(hdr)
          if h.startswith("smsp__pcsamp_warps_issue_stalled") and not h.endswith("not_issued") and r[i] not in ("", "n/a")]
    tot = sum(v for v, _ in st) or 1
    print("  stalls: " + ", ".join("%s %.0f%%" % (h.replace("smsp__pcsamp_warps_issue_stalled_", ""), 100 * v / tot)
                                   for v, h in sorted(st, reverse=True)[:6]))
