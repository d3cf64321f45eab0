"""Prints the metrics we track from an `ncu --page raw --csv` export.  Usage: ncu_summary.py raw.csv"""
import csv
import sys

WANT = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps',
        'smsp__issue_active.avg.pct', 'smsp__inst_executed.sum', 'sm__inst_executed_pipe_fp64.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'l1tex__data_bank_conflicts_pipe_lsu.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed_op_shared_atom.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum', 'smsp__inst_executed_op_shared_atom_dot_alu.sum',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic']


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if 'issue_stalled' in h and h.endswith('per_issue_active.ratio')]
    for r in rows[2:]:
        print('----')
        for w in WANT:
            if w in idx:
                print(f'{w} [{units[idx[w]]}] {r[idx[w]][:100]}')
        st = sorted(((float(r[idx[h]].replace(',', '') or 0), h) for h in stall), reverse=True)[:6]
        for v, h in st:
            print(f'  stall {h.split("issue_stalled_")[1].split("_per_issue")[0]}: {v:.2f}')


if __name__ == '__main__':
    main()
