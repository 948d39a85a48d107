"""Prints block 0's time line from an ACB200_POLICY_TRACE dump (development aid)."""
import sys
for fn in sys.argv[1:]:
    d = {}
    for ln in open(fn):
        i, t = ln.split(); d[int(i)] = int(t)
    m = [d[i] for i in sorted(d) if i < 512]; e = [d[i] for i in sorted(d) if i >= 1024]
    t0 = min(m[0], e[0])
    print(fn, "total cycles", max(m[-1], e[-1]) - t0)
    chunks = [1, 4, 4, 1, 4, 1, 1, 1]
    idx = 1
    for l, nc in enumerate(chunks):
        s = f"L{l}: act_ready@{m[idx]-t0:7d}"; idx += 1
        for c in range(nc):
            s += f" | c{c} start@{m[idx]-t0:7d} issued+{m[idx+1]-m[idx]:5d}"; idx += 2
        print(s)
    idx = 1; s = "EPI:"
    for k in range(17):
        s += f" [{e[idx]-t0}-{e[idx+1]-t0} d={e[idx+1]-e[idx]}]"; idx += 2
    print(s)
