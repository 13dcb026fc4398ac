"""Turns an `ncu --set full` report into the small CSV that is committed under profiles/ (and that bench.py reads for
`roofline.traffic`): one row per profiled launch, kernel name first, a header row of metric names and a row of units.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/r02_ncu_full_gemm_70b.csv [--match gemm]
"""
import csv
import subprocess
import sys

COLUMNS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
           "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size", "launch__block_size",
           "launch__cluster_dim_x", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    match = sys.argv[sys.argv.index("--match") + 1] if "--match" in sys.argv else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    head, units = rows[0], rows[1]
    kn = head.index("Kernel Name")
    cols = [(c, head.index(c)) for c in COLUMNS if c in head]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["Kernel Name"] + [c for c, _ in cols])
        w.writerow([""] + [units[i] for _, i in cols])
        for r in rows[2:]:
            if match in r[kn]:
                w.writerow([r[kn]] + [r[i] for _, i in cols])


if __name__ == "__main__":
    main()
