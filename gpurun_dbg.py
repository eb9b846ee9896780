import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
import golden_io
from dls_b200.batch import priorities, schedule_batch
from oracle import oracle
cases = golden_io.load("cfg1_example_and_edges.npz")
for c in cases:
    b = golden_io.batch_of([c])
    got = priorities(b); want = oracle.priorities(b)
    bad = [k for k in got if not np.array_equal(got[k].view(np.uint8), want[k].view(np.uint8))]
    r = schedule_batch(b, 8, peak=True, cached=True).host()
    ok = np.array_equal(r.node_of[0], c.exp[3]["node_of"])
    print(c.label, "n", c.n, "E", c.dag.E, "prio mismatch:", bad, "mru general ok:", ok)
    if bad:
        for k in bad: print("   ", k, got[k].tolist(), want[k].tolist())
