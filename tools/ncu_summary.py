"""Summarise an .ncu-rep (ncu --set full) into a small CSV for profiles/: one row per captured launch."""
import csv, subprocess, sys

KEYS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__shared_mem_per_block_dynamic", "dyn_smem"),
        ("sm__cycles_elapsed.avg.per_second", "sm_clock"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts")]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    with open(out, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel"] + [f"{n} [{units[col[k]]}]" if k in col else n for k, n in KEYS])
        for r in rows[2:]:
            name = r[col["Kernel Name"]]
            name = name.split("(")[0].replace("void ", "")
            w.writerow([name] + [r[col[k]] if k in col else "" for k, _ in KEYS])
    print(open(out).read())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
