"""Per-step timeline of the bichar kernel's roles from a CTCG_BICHAR_TRACE file.

For every step: the barrier release time, and per role the time its slowest warp
finished its work of the step (relative to the release that opened the step)."""
import struct
import sys

import numpy as np


def main(path, lo=None, hi=None):
    raw = open(path, 'rb').read()
    warps, steps, nmain, T = struct.unpack('4i', raw[:16])
    a = np.frombuffer(raw[16:], dtype=np.int64).reshape(warps, steps, 2)
    roles = {'main': range(0, nmain), 'prod': range(nmain, nmain + 4), 'reco': range(nmain + 4, nmain + 8),
             'scale': [nmain + 8], 'io': [nmain + 9]}
    used = np.where(a[0, :, 0] > 0)[0]
    n = used.max() + 1
    rel = a[:, :n, 1].max(axis=0)            # barrier release per step (any warp's stamp after the barrier)
    dur = np.diff(rel)
    print('steps', n, 'total cycles', rel[-1] - rel[0], 'mean step', dur.mean(), 'median', np.median(dur))
    # work end of step s+1 relative to release of step s
    out = {}
    for name, ws in roles.items():
        ws = list(ws)
        end = a[ws, 1:n, 0].max(axis=0) - rel[:-1]
        out[name] = end
    lo = lo or 0
    hi = hi or n - 1
    for name, end in out.items():
        seg = end[lo:hi]
        print('%-6s work-end after release: mean %7.1f  p50 %7.1f  p90 %7.1f  max %7.1f' % (
            name, seg.mean(), np.median(seg), np.percentile(seg, 90), seg.max()))
    seg = dur[lo:hi]
    print('step   duration:                mean %7.1f  p50 %7.1f  p90 %7.1f  max %7.1f' % (
        seg.mean(), np.median(seg), np.percentile(seg, 90), seg.max()))
    lag = np.argmax(np.stack([out[k][lo:hi] for k in out]), axis=0)
    names = list(out)
    print('slowest role per step:', {names[i]: int((lag == i).sum()) for i in range(len(names))})


if __name__ == '__main__':
    main(sys.argv[1], *(int(x) for x in sys.argv[2:4]))
