"""Summarise an `ncu --set full --import-source on` capture of the bichar kernel by CTA role.

    ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > X.csv
    python profiles/ncu_roles.py X.csv pytorch-asr_b200/csrc/kernels_bichar.cu

Prints executed warp-instructions and stall samples per role function / sweep, the
most-sampled SASS instructions (a barrier wait is charged to the instruction after
BAR.SYNC) and the opcode mix."""
import bisect
import collections
import csv
import re
import sys


def main(csv_path, src_path, top=25):
    rows = list(csv.reader(open(csv_path)))
    cur_file, line, seen = None, None, {}
    for r in rows:
        if not r:
            continue
        if r[0] == 'File Path':
            cur_file = r[1].split('/')[-1]
            continue
        if r[0] in ('Function Name', 'Line No'):
            continue
        if r[0] != '':
            line = (cur_file, int(r[0]))
            continue
        try:
            n, s = int(r[7]), int(r[6])
        except (ValueError, IndexError):
            continue
        a = r[2]
        if a in seen and seen[a][0][0] == src_path.split('/')[-1]:
            continue
        seen[a] = (line, n, s, r[3].strip())
    tot = sum(v[1] for v in seen.values())
    ts = sum(v[2] for v in seen.values())
    print('warp-instructions', tot, 'samples', ts)
    marks = []
    for i, l in enumerate(open(src_path).read().splitlines(), 1):
        m = re.match(r'__device__ __forceinline__ void (\w+)\(', l)
        if m:
            marks.append((i, m.group(1)))
        if 'between_sweeps(c)) return' in l and marks:
            marks.append((i, marks[-1][1].split(':')[0] + ':bwd'))
        if l.startswith('__global__'):
            marks.append((i, 'kernel'))
    marks.sort()
    agg, sm, ops = collections.Counter(), collections.Counter(), collections.Counter()
    for a, (ln, n, s, txt) in seen.items():
        op = txt.split()
        op = op[1] if op[0].startswith('@') else op[0]
        ops[op.split('.')[0]] += n
        if ln[0] != src_path.split('/')[-1]:
            name = 'inlined helpers (' + ln[0] + ')'
        else:
            k = bisect.bisect_right([m[0] for m in marks], ln[1]) - 1
            name = marks[k][1] if k >= 0 else 'pre'
        agg[name] += n
        sm[name] += s
    for k, v in agg.most_common():
        print('%-42s %12d %5.1f%%  samples %5.1f%%' % (k, v, 100 * v / tot, 100 * sm[k] / ts))
    print('--- most sampled instructions')
    for a, (ln, n, s, txt) in sorted(seen.items(), key=lambda kv: -kv[1][2])[:top]:
        print('%5.1f%%  %s:%d  %s  (x%d)' % (100 * s / ts, ln[0][:18], ln[1], txt[:64], n))
    print('--- opcode mix', [(o, round(100 * n / tot, 1)) for o, n in ops.most_common(14)])


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2])
