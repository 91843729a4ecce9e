"""Sums dram__bytes_read.sum + dram__bytes_write.sum per kernel family over ONE assemble from an ncu CSV log
(ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file x.csv <cmd>) and
records it in profiles/ncu_traffic.json under <family>_n<n>_w<world>, the key bench.py looks up.

python tools/ncu_traffic.py x.csv <n> <assembles captured> <capture id> [world]"""
import csv, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path, n, reps, cap = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
world = int(sys.argv[5]) if len(sys.argv) > 5 else 1
rows = [r for r in csv.reader(l for l in open(path) if not l.startswith('==')) if r]
hdr = rows[0]
ik, im, iv, iu = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
fam = {'pk': r'pk_kernel', 'ek': r'ek_kernel', 'rk': r'rk_kernel', 'fg': r'force_gather'}
tot = {k: 0.0 for k in fam}
tim = {k: 0.0 for k in fam}
scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 'second': 1e3, 's': 1e3}
for r in rows[1:]:
    for k, pat in fam.items():
        if re.search(pat, r[ik]):
            val = float(r[iv].replace(',', '')) * scale.get(r[iu], 1.0)
            if r[im].startswith('dram__bytes'):
                tot[k] += val
            elif r[im].startswith('gpu__time_duration'):
                tim[k] += val
out_path = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
try:
    out = json.load(open(out_path))
except Exception:
    out = {}
for k in fam:
    if tot[k] > 0:
        out[f"{k}_n{n}_w{world}"] = {"bytes": tot[k] / reps, "ms_under_ncu": tim[k] / reps, "capture": cap,
                                      "assembles_captured": reps}
json.dump(out, open(out_path, 'w'), indent=1, sort_keys=True)
print(json.dumps({k: v for k, v in out.items() if v.get("capture") == cap}, indent=1))
