#!/usr/bin/env python3
"""Condense an .ncu-rep (ncu --set full) into the few counters the roofline argument needs.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/<name>.txt
"""
import csv
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem/block"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit (smem), blocks"),
    ("launch__occupancy_limit_registers", "occupancy limit (regs), blocks"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % of peak"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("dram__bytes.sum.per_second", "DRAM throughput"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
    ("dram__bytes_write.sum.per_second", "DRAM write throughput"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 % of peak"),
    ("l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed", "SM->L2 request port busy %"),
    ("l1tex__m_l1tex2xbar_write_bytes.sum", "SM->L2 write bytes"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM read bytes"),
    ("sm__cycles_elapsed.avg", "SM cycles elapsed"),
    ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "L1 global-load sectors"),
    ("lts__t_sectors_srcunit_tex_op_read.sum", "L2 read sectors from SMs"),
    ("lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum", "  of which missed"),
    ("lts__t_sectors_srcunit_tex_op_write.sum", "L2 write sectors from SMs"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe (hmma) active %"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor instructions"),
    ("smsp__issue_active.avg.pct", "issue slots busy %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"# {rep}: ncu --set full --clock-control none (per launch; cold-ish caches, serialised)")
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        short = name.split("(")[0].split("::")[-1]
        print(f"\n== {short}   [{name[:110]}]")
        for key, label in WANT:
            if key in idx and r[idx[key]] not in ("", "n/a"):
                print(f"  {label:38s} {r[idx[key]]:>16s} {units[idx[key]]}")
        if "dram__bytes_read.sum" in idx:
            def gb(k):
                v, u = float(r[idx[k]]), units[idx[k]]
                return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
            print(f"  {'traffic (read+write)':38s} {(gb('dram__bytes_read.sum') + gb('dram__bytes_write.sum')) / 1e6:16.1f} MB")


if __name__ == "__main__":
    main()
