"""Summarises an .ncu-rep (ncu --set full) into a small CSV for profiles/ (run where ncu is installed)."""
import csv
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__bytes_read.sum.per_second',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'smsp__inst_executed.sum',
        'sm__cycles_active.min', 'sm__cycles_active.avg', 'sm__cycles_active.max', 'sm__cycles_elapsed.max',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__inst_executed_pipe_tensor.sum']


def short_name(k):
    """`void ssb::<unnamed>::kernel<...>(Params)` -> `kernel<...>` without namespaces."""
    import re
    k = re.sub(r'<unnamed>::|unnamed>::|ssb::|^void ', '', k)   # some ncu pages print `<unnamed>::` without its opening bracket
    return k.split('(')[0][:70]


def main(rep, out, title):
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index('Kernel Name')
    data = [r for r in data if 'at::' not in r[ki]]   # torch's own kernels (input generation) are not ours
    lines = ['# ' + title, '# kernels: ' + ' | '.join(sorted({r[ki][:100] for r in data})),
             'metric,unit,' + ','.join('launch_%d' % i for i in range(len(data))),
             'kernel,,' + ','.join('"%s"' % short_name(r[ki]) for r in data)]
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            lines.append('%s,%s,%s' % (w, units[i], ','.join(r[i].replace(',', '') for r in data)))
    for i, h in enumerate(hdr):
        if 'issue_stalled' in h and 'per_issue_active' in h:
            lines.append('%s,%s,%s' % (h, units[i], ','.join(r[i] for r in data)))
    open(out, 'w').write('\n'.join(lines) + '\n')
    print('\n'.join(lines[:14]))


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else sys.argv[1])
