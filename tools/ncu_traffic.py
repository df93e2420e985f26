#!/usr/bin/env python3
"""DRAM bytes per phase and launch shares from ncu CSV dumps.

  ncu_traffic.py traffic <raw.csv>      -> per-phase dram__bytes_read+write of one tools/profile_commit.py --reps 1 capture
  ncu_traffic.py shares <launches.csv>  -> per-kernel share of the summed gpu__time_duration of a launch list
"""
import csv
import collections
import json
import sys


def rows_raw(path):
    r = list(csv.reader(open(path)))
    h, units = r[0], r[1]
    return h, units, r[2:]


def to_bytes(v, unit):
    v = float(v.replace(',', ''))
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12}[unit]


def traffic(path):
    h, units, rows = rows_raw(path)
    ki, ri, wi = h.index('Kernel Name'), h.index('dram__bytes_read.sum'), h.index('dram__bytes_write.sum')
    out = collections.OrderedDict(ifft=0.0, lde=0.0, leaf_hash_kernel=0.0, tree=0.0)
    phase = None
    for row in rows:
        k = row[ki]
        b = to_bytes(row[ri], units[ri]) + to_bytes(row[wi], units[wi])
        if k.startswith('ntt_pass_hi') or 'ntt_pass_hi' in k:
            phase = 'ifft' if '<0, 0' in k else 'lde'
            out[phase] += b
        elif 'ntt_pass_lo' in k:
            out[phase] += b
        elif 'leaf_hash' in k:
            out['leaf_hash_kernel'] += b
        elif 'merkle' in k:
            out['tree'] += b
    print(json.dumps(out, indent=1))


def shares(path):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith('=='))]
    h = rows[0]
    ki, mi, vi = h.index('Kernel Name'), h.index('Metric Name'), h.index('Metric Value')
    ui = h.index('Metric Unit')
    tot = collections.Counter()
    cnt = collections.Counter()
    for r in rows[1:]:
        if len(r) <= vi or r[mi] != 'gpu__time_duration.sum':
            continue
        v = float(r[vi].replace(',', '')) * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(r[ui], 1e-6)
        name = r[ki].split('(')[0]
        tot[name] += v
        cnt[name] += 1
    s = sum(tot.values())
    print('total kernel time %.1f ms over %d launches' % (s, sum(cnt.values())))
    for k, v in tot.most_common():
        print('%-48s launches %5d  total %9.3f ms  share %5.1f %%' % (k[:48], cnt[k], v, 100 * v / s))


if __name__ == '__main__':
    {'traffic': traffic, 'shares': shares}[sys.argv[1]](sys.argv[2])
