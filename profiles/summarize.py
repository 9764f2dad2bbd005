"""Summarise an .ncu-rep (ncu --set full) into a small text file: python profiles/summarize.py rep.ncu-rep out.txt"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "sm__cycles_elapsed.max", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed"]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        for vals in rows[2:]:
            d = dict(zip(hdr, vals))
            f.write(f"== kernel: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}\n")
            for h, u, v in zip(hdr, units, vals):
                if any(h == k or h.startswith(k) for k in KEYS) and "per_second" not in h and "pct_of_peak_sustained_elapsed" not in h.replace("sm__throughput", ""):
                    f.write(f"{h} [{u}] = {v}\n")
            st = []
            for h, v in zip(hdr, vals):
                if "pcsamp_warps_issue_stalled" in h and "not_issued" not in h:
                    try:
                        st.append((float(v.replace(",", "")), h.split("issue_stalled_")[1]))
                    except ValueError:
                        pass
            tot = sum(x for x, _ in st) or 1.0
            f.write("-- warp issue-stall samples (share):\n")
            for x, h in sorted(st, reverse=True)[:10]:
                f.write(f"   {100 * x / tot:5.1f}%  {h}\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
