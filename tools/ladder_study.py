"""CPU study of ladder search sizes with the lane-emulated device code (node counts per item)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import go_oracle
from tests import hostsim
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
b, a, ko, pl = go_oracle.random_playouts(n, seed=int(sys.argv[2]) if len(sys.argv) > 2 else 77)
t = time.time()
out, st, items = hostsim.features49(b, a, ko, pl, item_cap=n * 40)
print("n=%d  %.1fs  items=%d nodes=%d  mean %.1f nodes/item" % (n, time.time() - t, st[3], st[2], st[2] / st[3]))
print("percentiles p50 %d p90 %d p99 %d p99.9 %d max %d (pos %d q %d kind %d maxdepth %d)" % (
    np.percentile(items, 50), np.percentile(items, 90), np.percentile(items, 99), np.percentile(items, 99.9), items.max(),
    st[5], st[6], st[7], st[8]))
print("top 10:", np.sort(items)[-10:].tolist(), " share of top 10 in total: %.2f" % (np.sort(items)[-10:].sum() / items.sum()))
if os.environ.get("CHECK"):
    want = go_oracle.features(b, a, ko, pl, ["ladder_capture", "ladder_escape"])
    print("parity vs oracle:", np.array_equal(out[:, 44:46], want))
