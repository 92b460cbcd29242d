"""Aggregate an ncu report's source page per CUDA source line:
   python profiles/ncu_source_lines.py <report.ncu-rep> [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No')
hdr = rows[hi]
ci = hdr.index('Instructions Executed'); si = hdr.index('# Samples')
bi = hdr.index('stall_barrier'); 
tot_i = tot_s = 0; lines = []
for r in rows[hi + 1:]:
    if len(r) <= ci or r[2] != '-':   # '-' address marks a source-line aggregate row
        continue
    try: n = int(r[ci]); sm = int(r[si]); sb = int(r[bi])
    except ValueError: continue
    tot_i += n; tot_s += sm
    lines.append((n, sm, sb, r[0], r[1].strip()[:110]))
print(f'total warp-instructions {tot_i}, samples {tot_s}')
for n, sm, sb, ln, src in sorted(lines, reverse=True)[:top]:
    print(f'{100*n/tot_i:5.1f}% inst {100*sm/max(1,tot_s):5.1f}% smp (barrier {sb:6d}) L{ln}: {src}')
