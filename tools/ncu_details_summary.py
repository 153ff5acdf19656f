"""Condense `ncu -i X.ncu-rep --page details --csv` (+ optionally `--page raw --csv`) into the per-kernel table kept under profiles/.
Usage: python tools/ncu_details_summary.py details.csv [raw.csv] > profiles/rN_kernels_full_summary.txt"""
import collections
import csv
import sys

WANT = ["Duration", "DRAM Throughput", "Memory Throughput", "Compute (SM) Throughput", "Executed Ipc Active", "Issue Slots Busy",
        "Executed Instructions", "Registers Per Thread", "Dynamic Shared Memory Per Block", "Theoretical Occupancy", "Achieved Occupancy",
        "L1/TEX Hit Rate", "L2 Hit Rate", "No Eligible", "Warp Cycles Per Issued Instruction", "Block Limit Registers", "Block Limit Shared Mem",
        "Cluster Size", "Grid Size", "Block Size"]
RAW = ["dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_bytes.sum",
       "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
       "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
       "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active", "gpu__time_duration.sum",
       "smsp__inst_executed.sum", "sm__cycles_elapsed.max"]


def main():
    rows = list(csv.DictReader(open(sys.argv[1])))
    per = collections.OrderedDict()
    for r in rows:
        per.setdefault((r["ID"], r["Kernel Name"], r["Grid Size"], r["Block Size"]), []).append(r)
    raw = {}
    if len(sys.argv) > 2:
        rr = list(csv.reader(open(sys.argv[2])))
        hdr = rr[0]
        for line in rr[2:]:
            d = dict(zip(hdr, line))
            raw[d["ID"]] = d
        units = dict(zip(hdr, rr[1]))
    for (kid, name, grid, block), items in per.items():
        print(f"== [{kid}] {name[:110]}\n   grid {grid} block {block}")
        seen = set()
        for r in items:
            if r["Metric Name"] in WANT and (r["Section Name"], r["Metric Name"]) not in seen:
                seen.add((r["Section Name"], r["Metric Name"]))
                print(f"   {r['Metric Name']:42s} {r['Metric Value']:>16s} {r['Metric Unit']:14s} ({r['Section Name']})")
        if kid in raw:
            for k in RAW:
                if k in raw[kid] and raw[kid][k] != "":
                    print(f"   raw {k:90s} {raw[kid][k]} {units.get(k, '')}")


if __name__ == "__main__":
    main()
