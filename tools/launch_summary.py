"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): per-kernel totals of one MSM pass in the file —
the FIRST large one (the headline's resident 2^20 step comes first in bench.py), or the last with --last.
    python tools/launch_summary.py gpurun_out/launches.csv [--detail] [--last]"""
import csv, re, sys
from collections import OrderedDict
rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
r = list(csv.reader(rows))
hdr = r[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
ks = [(re.sub(r'^void ', '', re.sub(r'[<(].*', '', x[ki])).split('::')[-1], float(x[vi].replace(',', '')) / (1000 if x[ui] == 'ns' else 1)) for x in r[1:]]
starts = [i for i, k in enumerate(ks) if k[0] == 'msm_digits' and (i == 0 or ks[i - 1][0] != 'msm_digits')]
def acc_us(i):
    return next((v for n, v in ks[i:i + 30] if n == 'msm_accumulate'), 0.0)
big = [i for i in starts if acc_us(i) >= 1000.0]
last = starts[-1] if '--last' in sys.argv or not big else big[0]
# the pass ends with msm_finish (multi-launch path) or with the last msm_tail / msm_tail_proj launch before another kernel family
end = last + 1
while end < len(ks) and ks[end][0].startswith(('msm_', 'scan_')) and ks[end][0] != 'msm_digits':
    end += 1
ks = ks[:end]
agg = OrderedDict(); tot = 0
for n, v in ks[last:]:
    a = agg.setdefault(n, [0, 0]); a[0] += v; a[1] += 1; tot += v
for n, (v, c) in agg.items():
    print(f"{n:28s} x{c:<3d} {v:9.1f} us  {100 * v / tot:5.1f} %")
print(f"{'total':28s}      {tot:9.1f} us")
if '--detail' in sys.argv:
    print('; '.join(f"{n} {v:.1f}" for n, v in ks[last:] if n.startswith(('msm_pair', 'msm_weighted', 'msm_group', 'msm_fold', 'msm_bucket'))))
