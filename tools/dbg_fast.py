import sys, os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orb_slam2_test_b200 import ORBextractor, synth
from oracle.pyoracle import Oracle
ext = ORBextractor(1000, 1.2, 8, 20, 7); ora = Oracle(1000)
img = synth.make_frame(synth.TUM_SHAPE, seed=3, frame=0)
ext(img); ora.extract(img)
for l in range(8):
    xs, ys, sc = ext.debug_candidates(0, l)
    c = ora.candidates(l)
    a = set(zip(xs.tolist(), ys.tolist(), sc.tolist())); b = set(zip(c["x"].tolist(), c["y"].tolist(), c["score"].tolist()))
    print("L", l, "gpu", len(a), "dups", len(xs) - len(a), "oracle", len(b), "missing", len(b - a), "extra", len(a - b))
    if l == 0:
        miss = sorted(b - a)
        from collections import Counter
        wc, hc = 31, 32
        cells = Counter(((y - 3) // hc, (x - 3) // wc) for x, y, s in miss)
        print("cells with misses", len(cells), list(cells.items())[:20])
        allc = Counter(((y - 3) // hc, (x - 3) // wc) for x, y, s in b)
        gc = Counter(((y - 3) // hc, (x - 3) // wc) for x, y, s in a)
        for k in list(cells)[:8]:
            print(k, "oracle", allc[k], "gpu", gc[k], [m for m in miss if ((m[1]-3)//hc, (m[0]-3)//wc) == k][:6])
            print("   gpu pts", sorted(p for p in a if ((p[1]-3)//hc, (p[0]-3)//wc) == k)[:10])
            print("   ora pts", sorted(p for p in b if ((p[1]-3)//hc, (p[0]-3)//wc) == k)[:10])
