"""Summarise an .ncu-rep (ncu --set full) into the text committed under profiles/.

    python tools/ncu_summary.py gpurun_out/r1_prof_chain.ncu-rep > profiles/r1_prof_chain_summary.txt

Per captured launch: duration, DRAM bytes (read/write/total), DRAM and SM throughput, occupancy
and its limiter, registers, shared memory, and the top warp-stall reasons (pc sampling).
"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm throughput % of peak"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("sm__warps_active.avg.per_cycle_active", "active warps / SM"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__waves_per_multiprocessor", "waves / SM"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_static", "static smem / block"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
    ("launch__occupancy_limit_registers", "occupancy limit (registers), blocks"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit (smem), blocks"),
    ("launch__occupancy_limit_warps", "occupancy limit (warps), blocks"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads / instruction"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
]


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in h]
    print("# %s  (ncu --set full --clock-control none; per-launch, cold-cache, serialised)" % path.split("/")[-1])
    for r in rows[2:]:
        print("\n== launch %s: %s" % (r[idx["ID"]], r[idx["Kernel Name"]]))
        for k, label in KEYS:
            if k in idx:
                print("   %-40s %s %s" % (label, r[idx[k]], units[idx[k]]))
        if "dram__bytes_read.sum" in idx:
            def mb(k):
                v, u = float(r[idx[k]]), units[idx[k]]
                return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
            print("   %-40s %.3f Mbyte" % ("dram traffic (read+write)", mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum")))
        tot = sum(float(r[idx[s]] or 0) for s in stall) or 1.0
        top = sorted(((float(r[idx[s]] or 0), s) for s in stall), reverse=True)[:6]
        print("   warp stalls (pc samples): " + ", ".join(
            "%s %.1f%%" % (s.replace("smsp__pcsamp_warps_issue_stalled_", ""), 100 * v / tot) for v, s in top))


if __name__ == "__main__":
    main(sys.argv[1])
