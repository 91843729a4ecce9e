"""Per-source-line summary of an `ncu --page source --csv --print-source cuda,sass` dump:
python tools/ncu_lines.py dump.csv [top]  -> samples, executed warp instructions, shared wavefronts per CUDA line."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out, cur_file, hdr = [], None, None
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
        continue
    if r[0] == 'Line No':
        hdr = r
        continue
    if hdr is None or r[0] in ('Function Name',):
        continue
    if r[0] != '' and r[2] == '-':  # a CUDA line with aggregated metrics
        def col(name):
            try:
                v = r[hdr.index(name)]
                return float(v) if v not in ('', '-') else 0.0
            except ValueError:
                return 0.0
        out.append((col('Warp Stall Sampling (All Samples)'), col('Instructions Executed'), col('L1 Wavefronts Shared'),
                    col('L1 Wavefronts Shared Ideal'), cur_file, r[0], r[1].strip()[:90]))
tot_s = sum(o[0] for o in out) or 1
tot_i = sum(o[1] for o in out) or 1
tot_w = sum(o[2] for o in out) or 1
print(f"total samples {tot_s:.0f} instructions {tot_i:.0f} shared wavefronts {tot_w:.0f}")
for o in sorted(out, reverse=True)[:top]:
    print(f"{100*o[0]/tot_s:5.1f}% smp {100*o[1]/tot_i:5.1f}% ins {100*o[2]/tot_w:5.1f}% wf (ideal {o[3]/max(o[2],1):.2f}) {o[4]}:{o[5]} {o[6]}")
