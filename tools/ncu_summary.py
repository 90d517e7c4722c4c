"""Print the metrics we track from an .ncu-rep (CPU-side; needs only the ncu CLI)."""
import csv, io, subprocess, sys

KEYS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
    'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.per_cycle_active',
    'smsp__inst_executed.sum', 'sm__inst_executed_pipe_fp64.sum', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_lsu.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
    'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed',
    'sm__cycles_elapsed.avg', 'sm__cycles_active.avg', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
    'lts__t_bytes.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
    'smsp__inst_executed.avg.per_cycle_active', 'sm__inst_executed_pipe_tensor.sum', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
    'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'local_load', 'smsp__cycles_active.avg',
]


def main(path, launch=0, extra=()):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for li, vals in enumerate(rows[2:]):
        if launch is not None and li != launch:
            continue
        rec = dict(zip(hdr, vals))
        print(f"--- launch {li}: {rec.get('Kernel Name','?')[:100]}")
        for h, u, v in zip(hdr, units, vals):
            stall = 'smsp__average_warps_issue_stalled' in h and h.endswith('_per_issue_active.ratio') and 'not_issued' not in h
            if h in KEYS or stall or any(e in h for e in extra):
                if stall:
                    try:
                        if float(v) < 0.15:
                            continue
                    except ValueError:
                        pass
                print(f"{h:96s} {v:>18s} {u}")


if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2] != 'all' else (None if len(sys.argv) > 2 else 0), tuple(sys.argv[3:]))
