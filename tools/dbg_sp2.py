import sys, numpy as np
sys.path.insert(0, '/root/repo')
import rawalign_b200 as ra
from rawalign_b200 import ANCHOR_DTYPE, CHAIN_DTYPE
rng = np.random.default_rng(5)
read = rng.normal(0, 1, 4000).astype(np.float32)
ref = rng.normal(0, 1, 4000).astype(np.float32)
ctxs = []
for legacy in (0, 1):
    opt = ra.ri_mapopt_init(); opt.reserved = 8 * legacy; opt.dtw_min_score = -1e9
    ctxs.append(ra.Context(0, opt))
def run(shapes):
    # one chain whose parts have the given (al, bl) shapes
    q = [0]; t = [0]
    for al, bl in shapes:
        q.append(q[-1] + al - 1); t.append(t[-1] + bl - 1)
    a = np.zeros(len(q), ANCHOR_DTYPE); a["query_position"] = q[::-1]; a["target_position"] = t[::-1]
    ch = np.zeros(1, CHAIN_DTYPE); ch["n_anchors"] = len(a)
    out = []
    for c in ctxs:
        s, _, i, u = c.align_chains_batch(read, ref, a, ch, min_score_mode=0, min_score=np.array([-1e10], np.float32))
        out.append((float(s[0]), i, u))
    return out
for shapes in [[(5, 5)], [(6, 5)], [(5, 6)], [(3, 3)], [(3, 2)], [(9, 7)], [(5, 5), (6, 5)], [(5, 5)] * 3, [(6, 5)] * 3, [(25, 22)], [(6, 5)] * 40, [(5, 5)] * 40]:
    o = run(shapes)
    print(shapes[:3], len(shapes), 'new', o[0], 'legacy', o[1], 'OK' if o[0] == o[1] else '<<<<')
