"""Summarise an .ncu-rep (read here, no GPU) into one row per kernel launch: duration, DRAM bytes,
throughput percentages, occupancy, registers, top stall reasons.  usage: ncu_summary.py rep [out.md]"""
import csv, io, subprocess, sys

WANT = {
    "gpu__time_duration.sum": "dur_us",
    "dram__bytes_read.sum": "dram_rd_MB",
    "dram__bytes_write.sum": "dram_wr_MB",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed": "l1_pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "occ_pct",
    "launch__registers_per_thread": "regs",
    "launch__occupancy_limit_registers": "lim_reg",
    "launch__occupancy_limit_shared_mem": "lim_smem",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "xu_pct",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "fma_pct",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active": "alu_pct",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "lsu_pct",
    "smsp__inst_executed.sum": "inst_M",
    "lts__t_sectors_op_red.sum": "l2_red_sect_M",
    "lts__t_sectors_op_atom.sum": "l2_atom_sect_M",
}
STALL_PREFIX = "smsp__average_warps_issue_stalled_"


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    lines = []
    for r in data:
        name = r[idx["Kernel Name"]].split("(")[0][-60:]
        d = {}
        for m, short in WANT.items():
            if m in idx:
                try:
                    v = float(r[idx[m]].replace(",", ""))
                except ValueError:
                    continue
                u = units[idx[m]]
                if short == "dur_us":
                    v = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u in ("ms", "msecond") else v)
                if short.endswith("_MB"):
                    v = {"byte": v / 1e6, "Kbyte": v / 1e3, "Mbyte": v, "Gbyte": v * 1e3}.get(u, v)
                if short.endswith("_M"):
                    v = v / 1e6
                d[short] = v
        stalls = []
        for h, i in idx.items():
            if h.startswith(STALL_PREFIX) and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
                try:
                    stalls.append((float(r[i].replace(",", "")), h[len(STALL_PREFIX):].replace("_per_issue_active.ratio", "")))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        d["stalls"] = " ".join(f"{n}:{v:.1f}" for v, n in stalls[:4])
        d["grid"] = r[idx["Grid Size"]] if "Grid Size" in idx else ""
        d["block"] = r[idx["Block Size"]] if "Block Size" in idx else ""
        lines.append((name, d))
    cols = ["dur_us", "dram_rd_MB", "dram_wr_MB", "dram_pct", "l2_pct", "l1_pct", "sm_pct", "issue_pct", "fma_pct", "alu_pct",
            "xu_pct", "lsu_pct", "occ_pct", "regs", "inst_M", "l2_red_sect_M"]
    txt = ["| kernel | grid | block | " + " | ".join(cols) + " | top stalls (warps stalled per issue) |", "|" + "---|" * (len(cols) + 4)]
    for name, d in lines:
        txt.append(f"| {name} | {d['grid']} | {d['block']} | " + " | ".join(
            (f"{d[c]:.1f}" if c in d else "-") for c in cols) + f" | {d['stalls']} |")
    s = "\n".join(txt)
    print(s)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(s + "\n")


if __name__ == "__main__":
    main()
