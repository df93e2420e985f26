#!/usr/bin/env python3
"""Static instruction cost of one Poseidon permutation inside a compiled kernel.

Reads the SASS of `leaf_hash_kernel` (or --kernel) from the built object with cuobjdump, finds the two innermost
loops (full-round body x 8, partial-round body x 22) and the loop that encloses them (one permutation), and
counts per permutation: N = instructions, W = IMAD.WIDE / IMAD.HI, F = other FMA-heavy-pipe instructions,
A = ALU-pipe instructions.  Measured on B200: cycles per warp-permutation = 1.14 x (N + 2 W) for every variant
of the permutation tried this round, i.e. the kernel is bound by issue slots with a wide multiply worth three.

  python tools/sass_cost.py [--obj 0-kno-blobstreamx_b200/csrc/hash_kernels.o] [--json out.json]
"""
import argparse
import collections
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ALU = ('IADD3', 'LOP3', 'SHF', 'SEL', 'ISETP', 'MOV', 'LEA', 'PRMT', 'VIADD', 'CS2R', 'IABS')


def pipe(op):
    if op.startswith('IMAD.WIDE') or op.startswith('IMAD.HI'):
        return 'W'
    if op.startswith('IMAD'):
        return 'F'
    if op.startswith(ALU):
        return 'A'
    return 'O'


def analyse(obj, kernel):
    names = subprocess.check_output(['cuobjdump', '-sass', obj]).decode()
    fn = [m for m in re.findall(r'Function : (\S+)', names) if kernel in m]
    if not fn:
        raise SystemExit('kernel %s not found in %s' % (kernel, obj))
    sass = subprocess.check_output(['cuobjdump', '-sass', '-fun', fn[0], obj]).decode()
    ins = []
    for line in sass.splitlines():
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)(.*?);', line)
        if m:
            ins.append((int(m.group(1), 16), m.group(3), m.group(4)))
    loops = []
    for ad, op, rest in ins:
        if op.startswith('BRA'):
            m = re.search(r'0x([0-9a-f]+)', rest)
            if m and int(m.group(1), 16) < ad:
                loops.append((int(m.group(1), 16), ad))
    inner = sorted(l for l in loops if not any(o != l and l[0] <= o[0] and o[1] <= l[1] for o in loops))
    inner = sorted(sorted(inner, key=lambda l: l[0] - l[1])[:2])  # the two largest innermost loops: full, partial
    if len(inner) != 2:
        raise SystemExit('expected a full-round and a partial-round loop, found %r' % (inner,))
    enclosing = [o for o in loops if o not in inner and all(o[0] <= i[0] and i[1] <= o[1] for i in inner)]
    perm = max(enclosing, key=lambda o: o[1] - o[0])  # the loop that runs once per permutation (column chunks)

    def count(a, b, excl=()):
        c = collections.Counter()
        for ad, op, _ in ins:
            if a <= ad <= b and not any(x[0] <= ad <= x[1] for x in excl):
                c[pipe(op)] += 1
        return c
    tot = collections.Counter()
    parts = {}
    for (a, b), mult, name in zip(inner, (8, 22), ('full_round', 'partial_round')):
        c = count(a, b)
        parts[name] = {'instructions': sum(c.values()), 'wide': c['W'], 'fma_heavy_other': c['F'], 'alu': c['A'],
                       'other': c['O'], 'trips': mult}
        for k, v in c.items():
            tot[k] += v * mult
    c = count(perm[0], perm[1], excl=inner)
    parts['per_permutation_rest'] = {'instructions': sum(c.values()), 'wide': c['W']}
    for k, v in c.items():
        tot[k] += v
    n = sum(tot.values())
    return {'kernel': fn[0], 'object': os.path.relpath(obj, ROOT), 'static_instructions': len(ins),
            'per_permutation': {'instructions': n, 'wide': tot['W'], 'fma_heavy_other': tot['F'], 'alu': tot['A'],
                                'other': tot['O'], 'issue_slots_n_plus_2w': n + 2 * tot['W']},
            'bodies': parts}


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--obj', default=os.path.join(ROOT, '0-kno-blobstreamx_b200', 'csrc', 'hash_kernels.o'))
    ap.add_argument('--kernel', default='leaf_hash_kernel')
    ap.add_argument('--json')
    a = ap.parse_args()
    r = analyse(a.obj, a.kernel)
    print(json.dumps(r, indent=1))
    if a.json:
        json.dump(r, open(a.json, 'w'), indent=1)
