"""Summarise an .ncu-rep (read here with `ncu -i`, no GPU needed) into one line per profiled launch.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_xxx_ncu_summary.txt
"""
import csv
import io
import subprocess
import sys

WANT = [
    ('gpu__time_duration.sum', 'dur'),
    ('dram__bytes_read.sum', 'dram_rd'),
    ('dram__bytes_write.sum', 'dram_wr'),
    ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
    ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor_pipe%'),
    ('sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active', 'hmma%'),
    ('sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active', 'hmma_inst%'),
    ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2%'),
    ('l1tex__m_xbar2l1tex_read_bytes.sum', 'l2->sm'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps%'),
    ('launch__registers_per_thread', 'regs'),
    ('launch__grid_size', 'grid'),
    ('launch__cluster_size', 'cluster'),
]
rep = sys.argv[1]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    name = r[hdr.index('Kernel Name')].split('(')[0].replace('void ', '')
    parts = ['%-34s' % name[:34]]
    for key, short in WANT:
        if key in hdr:
            i = hdr.index(key)
            v = r[i]
            try:
                v = '%.4g' % float(v.replace(',', ''))
            except ValueError:
                pass
            parts.append('%s=%s%s' % (short, v, units[i] if units[i] not in ('', '%') else ''))
    print(' | '.join(parts))
