"""Per-step timeline of the bichar kernel's roles from a CTCG_BICHAR_TRACE file.

Stamp 0 of a step = the warp reached the step's barrier (its work is done), stamp 1 = first
instruction after the barrier was issued (NOT the release: BAR.SYNC blocks at the next
dependent instruction), stamps 2.. = intermediate points of the main warps' forward step."""
import struct
import sys

import numpy as np


def main(path, lo=None, hi=None):
    raw = open(path, 'rb').read()
    warps, steps, nmain, trs = struct.unpack('4i', raw[:16])
    a = np.frombuffer(raw[16:], dtype=np.int64).reshape(warps, steps, trs)
    roles = {'main': range(0, nmain), 'aux0-3': range(nmain, nmain + 4), 'aux4-7': range(nmain + 4, nmain + 8),
             'scale': [nmain + 8], 'io': [nmain + 9]}
    n = np.where(a[0, :, 0] > 0)[0].max() + 1
    arr = a[:, :n, 0]
    last = arr.max(axis=0)                    # barrier completes when the last warp arrives
    lo = lo or 1
    hi = hi or n - 1
    dur = np.diff(last)
    print('steps', n, 'cycles', last[-1] - last[0], 'step mean %.0f p50 %.0f' % (dur[lo:hi].mean(), np.median(dur[lo:hi])))
    for name, ws in roles.items():
        ws = list(ws)
        work = arr[ws, 1:] - last[None, :-1]  # arrival at barrier s+1 minus completion of barrier s
        seg = work[:, lo:hi]
        print('%-7s arrival after previous barrier completed: mean %7.1f  p50 %7.1f  max-warp mean %7.1f' % (
            name, seg.mean(), np.median(seg), seg.max(axis=0).mean()))
    if trs > 2:
        w = 1  # a main warp with 4 majors
        for slot in range(2, trs):
            st = a[w, 1:n, slot]
            ok = st[lo:hi] > 0
            if ok.any():
                d = (st[lo:hi] - last[:-1][lo:hi])[ok]
                print('main warp %d stamp %d: mean %7.1f p50 %7.1f after the previous barrier completed' % (w, slot, d.mean(), np.median(d)))


if __name__ == '__main__':
    main(sys.argv[1], *(int(x) for x in sys.argv[2:4]))
