"""Aggregates an `ncu --page source --csv --print-source cuda,sass` dump of pk_kernel by kernel phase (markers in
pk_kernel.cuh): share of stall samples, of executed instructions and the top stall reasons per phase.
python tools/ncu_regions.py dump.csv"""
import csv, collections, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = list(csv.reader(open(sys.argv[1])))
src = open(os.path.join(ROOT, "total-lagrangian-fea_b200/csrc/pk_kernel.cuh")).read().split('\n')
MARKS = [('Gauss-point step, Mooney', 'gp_mr'), ('Gauss-point step, SVK', 'gp_svk'), ('element forces: sum', 'force'),
         ('--- block step', 'block'), ('finished in place', 'finish'), ('ordered accumulation into', 'rmw'),
         ('--- write-out', 'wo_own(a)'), ('the lists for after the barrier', 'wo_stage'), ('pairs held by several warps', 'merge(b)'),
         ('entries of the merged blocks, all threads', 'wo_dep(c)'), ('Write-out of one entry segment', 'wo_segment_fn'),
         ('template <int W, bool DAMP, bool TANGENT, int MODEL>', 'kernel_head'),
         ('nobody reads the accumulators', 'tail(zero,desc)'), ("The warp's element pairs form ONE stream", 'stream')]
marks = [(i, n) for i, t in enumerate(src, 1) for k, n in MARKS if k in t]
def reg_of(f, l):
    if 'pk_kernel' not in f:
        return 'helpers'
    r = 'head(prefetch fns)'
    for m, n in marks:
        if int(l) >= m:
            r = n
    return r
hdr = cur = curfile = None
reg = collections.defaultdict(collections.Counter)
ins = collections.Counter()
wav = collections.Counter()
wid = collections.Counter()
gsec = collections.Counter()
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        curfile = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
    elif hdr is None or r[0] == 'Function Name':
        continue
    elif r[0] != '' and r[2] == '-':
        cur = reg_of(curfile, r[0])
    elif r[0] == '' and len(r) > 10:
        for i, h in enumerate(hdr):
            if h.startswith('stall_') and '(Not Issued)' not in h:
                try:
                    reg[cur][h] += float(r[i] or 0)
                except ValueError:
                    pass
        for name, ctr in (('Instructions Executed', ins), ('L1 Wavefronts Shared', wav), ('L1 Wavefronts Shared Ideal', wid),
                          ('L2 Theoretical Sectors Global', gsec)):
            try:
                ctr[cur] += float(r[hdr.index(name)] or 0)
            except ValueError:
                pass
tot = sum(sum(c.values()) for c in reg.values())
ti = sum(ins.values())
tw = sum(wav.values())
print(f"samples {tot:.0f} warp instructions {ti:.0f} shared wavefronts {tw:.0f} (ideal {sum(wid.values()):.0f}) global sectors {sum(gsec.values()):.0f}")
for k, c in sorted(reg.items(), key=lambda kv: -sum(kv[1].values())):
    s = sum(c.values())
    top = ', '.join(f"{h[6:]} {100*v/s:.0f}%" for h, v in c.most_common(4))
    print(f"{k:20s} samples {100*s/tot:5.1f}%  instr {100*ins[k]/ti:5.1f}%  smem wavefronts {100*wav[k]/max(tw,1):5.1f}% (x{wav[k]/max(wid[k],1):.2f} of ideal)  gsectors {gsec[k]:.3g} | {top}")
