"""Summarise an .ncu-rep (ncu --set full) into a small text table for profiles/.
Usage: python tools/ncu_summary.py report.ncu-rep [--json profiles/xxx.json] > profiles/xxx.txt
The JSON form (per launch: kernel, ms, DRAM bytes read / written, tensor-pipe %) is what bench.py reads
its `roofline.traffic` from."""
import csv
import io
import subprocess
import sys

METRICS = [
    'gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'launch__registers_per_thread', 'launch__grid_size',
    'launch__block_size', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
    'sm__pipe_tensor_cycles_active.max.pct_of_peak_sustained_elapsed',
    'sm__ops_path_tensor_op_utchmma_src_tf32_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed',
    'l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
    'l1tex__m_xbar2l1tex_read_bytes.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
    'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
    'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed',
    'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
    'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    name_i = hdr.index('Kernel Name')
    print('# %s' % rep)
    print('# kernels: %s' % ', '.join(sorted({r[name_i].split('(')[0] for r in rows[2:]})))
    for m in METRICS:
        if m in hdr:
            i = hdr.index(m)
            print('%-100s %-10s %s' % (m, units[i], '  '.join(r[i] for r in rows[2:])))
    if '--json' in sys.argv:
        import json

        def col(name, r, scale_units=True):
            if name not in hdr:
                return None
            i = hdr.index(name)
            v = float(r[i].replace(',', ''))
            u = units[i].lower()
            if scale_units:
                v *= {'gbyte': 1e9, 'mbyte': 1e6, 'kbyte': 1e3, 'byte': 1.0, 'ms': 1.0, 'us': 1e-3, 'ns': 1e-6,
                      'msecond': 1.0, 'usecond': 1e-3, 'nsecond': 1e-6, 'second': 1e3}.get(u, 1.0)
            return v
        out = []
        for r in rows[2:]:
            out.append({'kernel': r[name_i].split('(')[0].replace('void ', '').replace('kfb::<unnamed>::', ''),
                        'ms': col('gpu__time_duration.sum', r),
                        'dram_read_bytes': col('dram__bytes_read.sum', r), 'dram_write_bytes': col('dram__bytes_write.sum', r),
                        'tensor_pipe_pct': col('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', r, False),
                        'dram_pct': col('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', r, False)})
        with open(sys.argv[sys.argv.index('--json') + 1], 'w') as f:
            json.dump({'report': rep, 'launches': out}, f, indent=1)


if __name__ == '__main__':
    main()
