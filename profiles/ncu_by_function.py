"""Aggregate an ncu source page by device function (line ranges parsed from the .cuh)."""
import csv, re, subprocess, sys
rep, src = sys.argv[1], sys.argv[2]
text = open(src).read().split('\n')
starts = []
for i, l in enumerate(text, 1):
    m = re.match(r'^(?:template <[^>]*>\s*)?(?:__device__|__global__).*?\b(\w+)\s*\(', l)
    if m and not l.startswith(' '): starts.append((i, m.group(1)))
    elif re.match(r'^__global__|^__device__', l):
        m2 = re.search(r'(\w+)\s*\(', l)
        if m2: starts.append((i, m2.group(1)))
# also template<> lines followed by __global__/__device__
fn_at = {}
cur = 'header'
si = 0
starts.sort()
def fn_of(line):
    name = 'prologue'
    for i, n in starts:
        if i <= line: name = n
        else: break
    return name
out = subprocess.run(['ncu','-i',rep,'--page','source','--print-source','cuda,sass','--csv'],capture_output=True,text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i,r in enumerate(rows) if r and r[0]=='Line No')
hdr = rows[hi]; ci = hdr.index('Instructions Executed'); sm = hdr.index('# Samples'); bi = hdr.index('stall_barrier')
agg = {}
for r in rows[hi+1:]:
    if len(r) <= ci or r[2] != '-': continue
    try: ln=int(r[0]); n=int(r[ci]); s=int(r[sm]); b=int(r[bi])
    except ValueError: continue
    f = fn_of(ln); a = agg.setdefault(f,[0,0,0]); a[0]+=n; a[1]+=s; a[2]+=b
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
for f,a in sorted(agg.items(), key=lambda x:-x[1][0]):
    print(f'{f:22s} inst {100*a[0]/ti:5.1f}%  samples {100*a[1]/ts:5.1f}%  barrier-samples {100*a[2]/ts:5.1f}%')
