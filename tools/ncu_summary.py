"""Summarises an `ncu --page raw --csv` export: one block of selected metrics per kernel launch.
    ncu -i X.ncu-rep --page raw --csv > X_raw.csv ; python tools/ncu_summary.py X_raw.csv [--csv]"""
import csv
import sys

WANT = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.sum', 'smsp__inst_executed.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'smsp__warp_issue_stalled_barrier_per_warp_active.pct', 'smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct',
        'smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct', 'smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct',
        'smsp__warp_issue_stalled_wait_per_warp_active.pct', 'smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct',
        'smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct', 'smsp__warp_issue_stalled_no_instruction_per_warp_active.pct',
        'smsp__warp_issue_stalled_imc_miss_per_warp_active.pct', 'lts__t_bytes.sum', 'local_load', 'smsp__inst_executed_op_local_ld.sum',
        'smsp__inst_executed_op_local_st.sum']


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    idx = [(w, hdr.index(w)) for w in WANT if w in hdr]
    if '--grep' in sys.argv:      # every metric whose name contains one of the comma-separated patterns, one kernel per block
        pats = sys.argv[sys.argv.index('--grep') + 1].split(',')
        ki = hdr.index('Kernel Name')
        for r in rows[2:]:
            print('-----', r[ki][:90])
            for i, h in enumerate(hdr):
                if any(p in h for p in pats):
                    print(f"{h:95s} {r[i][:40]} {units[i]}")
        return
    if '--csv' in sys.argv:
        w = csv.writer(sys.stdout)
        w.writerow([a for a, _ in idx]); w.writerow([units[i] for _, i in idx])
        for r in rows[2:]:
            w.writerow([r[i] for _, i in idx])
        return
    for r in rows[2:]:
        print('-----')
        for name, i in idx:
            print(f"{name:75s} {r[i][:70]} {units[i]}")


if __name__ == '__main__':
    main()
