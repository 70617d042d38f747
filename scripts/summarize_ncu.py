"""ncu report -> small text summary for profiles/ (per kernel: duration, DRAM bytes, throughput, occupancy, stalls)."""
import csv, io, subprocess, sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_red.sum"]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out, "w") as f:
        f.write(f"# ncu --set full --clock-control none summary of {rep.split('/')[-1]}\n")
        for r in rows[2:]:
            f.write(f"\n## {r[idx['Kernel Name']]}\n")
            for k in KEYS:
                if k in idx:
                    f.write(f"{k} = {r[idx[k]]} {units[idx[k]]}\n")
            st = []
            for h, i in idx.items():
                if "issue_stalled" in h and h.endswith("ratio"):
                    try:
                        st.append((float(r[i]), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
                    except ValueError:
                        pass
            st.sort(reverse=True)
            f.write("top stalls (cycles per issue): " + ", ".join(f"{n}={v:.2f}" for v, n in st[:5]) + "\n")
    print("wrote", out)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
