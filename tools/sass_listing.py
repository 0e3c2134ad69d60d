#!/usr/bin/env python
"""SASS evidence for the sm_100a kernels: per-kernel mnemonic counts (tcgen05 = UTCHMMA / UTCBAR / LDTM / UTCATOMSWS,
TMA = UTMALDG / UBLKCP, mbarrier = SYNCS.*, cluster = UCGABAR_*, DSMEM = ST*.E.*CLUSTER / STAS) plus the instruction
listing of the tensor-core main loops.  Runs here (no GPU needed): cuobjdump on the in-tree libpctts.so.

    python tools/sass_listing.py profiles/r01_sass
writes  <prefix>_summary.txt  (every kernel, one line of counts)  and  <prefix>_tc_excerpt.txt  (the lines around each
UTCHMMA / UTMALDG / LDTM of the tcgen05 kernels).
"""
import collections
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(HERE, 'percivaltts_b200', 'libpctts.so')
KEYS = ['UTCHMMA', 'UTCBAR', 'UTCATOMSWS', 'LDTM', 'UTMALDG', 'UTMACCTL', 'UBLKCP', 'SYNCS', 'ELECT', 'UCGABAR', 'STAS', 'SHFL', 'FFMA',
        'HFMA2', 'LDG', 'STG', 'LDS', 'STS', 'RED', 'ATOM', 'MUFU', 'BAR.SYNC']


def demangle(names):
    out = subprocess.run(['c++filt'], input='\n'.join(names), capture_output=True, text=True).stdout.split('\n')
    return dict(zip(names, out))


def main():
    prefix = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, 'profiles', 'sass')
    sass = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True, check=True).stdout.split('\n')
    funcs = collections.OrderedDict()
    cur = None
    for ln in sass:
        m = re.search(r'Function : (\S+)', ln)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur is not None and re.search(r'/\*[0-9a-f]{4}\*/', ln):
            funcs[cur].append(ln.rstrip())
    names = demangle(list(funcs))
    with open(prefix + '_summary.txt', 'w') as f:
        f.write('# cuobjdump -sass percivaltts_b200/libpctts.so  (nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo)\n')
        f.write('# per kernel: number of SASS instructions and counts of selected mnemonics (zero counts omitted)\n')
        for k, lines in funcs.items():
            cnt = collections.Counter()
            for ln in lines:
                body = ln.split('*/', 1)[1] if '*/' in ln else ln
                for key in KEYS:
                    if re.search(r'\b' + re.escape(key), body):
                        cnt[key] += 1
            short = names[k].replace('(anonymous namespace)::', '')
            short = re.sub(r'^void ', '', short)
            short = re.sub(r'\((?!anonymous).*', '', short)
            f.write('%-46s insts=%5d  %s\n' % (short[:46], len(lines), ' '.join('%s=%d' % (a, cnt[a]) for a in KEYS if cnt[a])))
    with open(prefix + '_tc_excerpt.txt', 'w') as f:
        f.write('# SASS around the tcgen05 / TMA instructions of the tensor-core kernels (2 lines of context)\n')
        for k, lines in funcs.items():
            hits = [i for i, ln in enumerate(lines) if re.search(r'UTCHMMA|UTMALDG|LDTM|UTCBAR|UTCATOMSWS', ln)]
            if not hits:
                continue
            f.write('\n==== %s\n' % names[k][:200])
            keep = set()
            for i in hits:
                keep.update(range(max(0, i - 2), min(len(lines), i + 3)))
            last = -2
            for i in sorted(keep):
                if i != last + 1:
                    f.write('        ...\n')
                f.write(lines[i] + '\n')
                last = i
    print('wrote', prefix + '_summary.txt', prefix + '_tc_excerpt.txt')


if __name__ == '__main__':
    main()
