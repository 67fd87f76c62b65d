"""Development aid: summarise a BAGPU_CHOL_TRACE dump (task table + 4 globaltimer stamps per task)."""
import sys
import numpy as np
raw = open(sys.argv[1], "rb").read()
n = int(np.frombuffer(raw[:4], np.int32)[0])
tasks = np.frombuffer(raw[4:4 + 32 * n], np.int32).reshape(n, 8)
st = np.frombuffer(raw[4 + 32 * n:], np.uint64).reshape(n, 4).astype(np.float64)
t0 = st[:, 0].min()
st = (st - t0) / 1e3   # us
names = ["diag", "off", "backtile", "backrow", "backchain"]
print("tasks", n, "span %.1f us" % (st[:, 3].max()))
for k in range(5):
    m = tasks[:, 0] == k
    if not m.any():
        continue
    d = st[m]
    print("%-8s n=%6d  claim->ready %7.2f  ready->computed %7.2f  computed->published %7.2f  (mean us)" % (
        names[k], m.sum(), (d[:, 1] - d[:, 0]).mean(), (d[:, 2] - d[:, 1]).mean(), (d[:, 3] - d[:, 2]).mean()))
dg = st[tasks[:, 0] == 0]
print("phase 1 end %.1f us; diag-to-diag publish interval mean %.2f us" % (dg[:, 3].max(), np.diff(dg[:, 3]).mean()))
print("diag: ready->computed (factor) median %.2f ; claim->ready median %.2f" % (np.median(dg[:, 2] - dg[:, 1]), np.median(dg[:, 1] - dg[:, 0])))
br = st[tasks[:, 0] >= 2]
ch = tasks[tasks[:, 0] == 4]
print("phase 2: %.1f us from the first claim, %.1f us after phase 1; %d chains, longest %d rows" % (
    br[:, 3].max() - br[:, 0].min(), br[:, 3].max() - dg[:, 3].max(), len(ch), (ch[:, 1] - ch[:, 2] + 1).max() if len(ch) else 0))
# a few consecutive rows in detail
off = tasks[:, 0] == 1
for k in range(100, 103):
    d = st[(tasks[:, 0] == 0) & (tasks[:, 1] == k)][0]
    print("row %d diag: claim %.1f ready %.1f computed %.1f published %.1f" % (k, *d))
    for row in np.where(off & (tasks[:, 1] == k))[0][:3]:
        print("   off (%d,%d): claim %.1f accum-done %.1f diag-seen %.1f published %.1f terms %d" % (k, tasks[row, 2], *st[row], tasks[row, 5] - tasks[row, 4]))
