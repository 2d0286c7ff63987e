"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line:
samples, warp instructions, shared wavefronts, top stall reasons.  python tools/ncu_lines.py src.csv [top]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = cur_fn = hdr = None
agg = collections.defaultdict(lambda: [0, 0, 0, '', collections.Counter()])
STALLS = ('stall_barrier', 'stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_lg', 'stall_wait', 'stall_math',
          'stall_not_selected', 'stall_branch_resolving', 'stall_membar', 'stall_tex', 'stall_dispatch', 'stall_no_inst',
          'stall_selected', 'stall_sleep')
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        cur_fn = r[1].split('(')[0].split('::')[-1][:24]; continue
    if r[0] == 'Line No':
        hdr = r
        ie = hdr.index('Instructions Executed'); iw = hdr.index('L1 Wavefronts Shared') if 'L1 Wavefronts Shared' in hdr else ie; ism = hdr.index('# Samples')
        continue
    if hdr is None or r[0] == '':
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    k = (cur_fn, cur_file, ln)
    g = lambda i: int(r[i]) if r[i].isdigit() else 0
    agg[k][0] += g(ism); agg[k][1] += g(ie); agg[k][2] += g(iw); agg[k][3] = r[1].strip()[:74]
    for nm in STALLS:
        if nm in hdr:
            agg[k][4][nm[6:]] += g(hdr.index(nm))
for fn in sorted({k[0] for k in agg}):
    items = [(k, v) for k, v in agg.items() if k[0] == fn]
    ts = sum(v[0] for k, v in items); ti = sum(v[1] for k, v in items); tw = sum(v[2] for k, v in items)
    print(f"===== {fn}: samples {ts}, warp instr {ti}, shared wavefronts {tw}")
    for k, v in sorted(sorted(items, key=lambda kv: -kv[1][0])[:top], key=lambda kv: (kv[0][1], kv[0][2])):
        st = ', '.join(f"{a}={b}" for a, b in v[4].most_common(2))
        print(f"{k[1][:15]:15s}:{k[2]:4d} smp {v[0]:5d} {100*v[0]/max(ts,1):4.1f}% inst {v[1]:9d} wf {v[2]:9d} [{st}] {v[3]}")
