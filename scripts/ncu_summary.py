"""Text summary of an .ncu-rep (ncu -i ... --page raw --csv) with the metrics the bench line and DESIGN.md quote.

  python scripts/ncu_summary.py gpurun_out/prof_mma.ncu-rep > profiles/rNN_mma_loop_ncu.txt
"""
import csv
import io
import re
import subprocess
import sys

PAT = re.compile(r'^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum(\.per_second)?|gpu__dram_throughput\.avg\.pct|'
                 r'lts__throughput\.avg\.pct|lts__t_bytes\.sum$|sm__throughput\.avg\.pct|sm__warps_active\.avg\.pct|'
                 r'sm__pipe_tensor_cycles_active\.avg\.pct|sm__pipe_fma_cycles_active\.avg\.pct|sm__pipe_alu_cycles_active\.avg\.pct|'
                 r'sm__inst_executed_pipe_(lsu|xu|uniform|tensor_subpipe_hmma)\.avg\.pct|l1tex__data_pipe_lsu_wavefronts(_mem_shared)?\.sum(\.pct.*)?$|'
                 r'smsp__inst_executed\.sum$|smsp__cycles_active\.avg$|launch__(registers_per_thread$|grid_size|block_size|shared_mem_per_block_dynamic|'
                 r'occupancy_limit.*)|smsp__average_warps?_issue_stalled_[a-z_]+_per_issue_active\.ratio|sm__cycles_elapsed\.max$|'
                 r'smsp__warp_issue_stalled_[a-z_]+_per_warp_active\.pct)')
rep = sys.argv[1]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
print(f'# {rep}: ncu --set full --clock-control none (cold cache, serialised replays: compare shares, not absolutes)')
for r in rows[2:]:
    print('\n== ' + r[hdr.index('Kernel Name')] + '   grid ' + r[hdr.index('Grid Size')] + ' block ' + r[hdr.index('Block Size')])
    for h, u, v in zip(hdr, units, r):
        if PAT.match(h) and v not in ('', 'n/a'):
            print(f'   {h:<88} {v} {u}')
