"""Checks the per-row closed form of the reference's banded cell set (dtw.cpp:273-520) against the
anti-diagonal enumeration.  Dev tool; the same check lives in tests/."""
import random, sys

def cellset(n, m, R):
    extra = ((n - m) * R + n - 1) // n
    Rp = R + extra
    P = Rp + (1 if Rp % 2 == 0 else 0); S = Rp + (1 if Rp % 2 == 1 else 0)
    cells = set()
    def seg(si, sj, ln):
        for o in range(ln):
            i, j = si - o, sj + o
            if 0 <= i < n and 0 <= j < m: cells.add((i, j))
    cells.add((0, 0)); r = 0
    for c in range(1, n):
        if (r + 1) * n <= m * c:
            r += 1; seg(c + S // 2 - 1, r - S // 2, S)
        seg(c + P // 2, r - P // 2, P)
    return cells, Rp, P, S

def rows_formula(n, m, Rp):
    odd = Rp % 2; h = (Rp - 1) // 2 if odd else Rp // 2
    out = []
    for i in range(n):
        ih = i + h + odd; il = i - h - odd
        jhi = m - 1 if ih > n - 1 else min(m - 1, (m * ih) // n + h)
        jlo = 0 if il < 0 else max(0, (m * il) // n - h)
        out.append((jlo, jhi))
    return out

random.seed(1); bad = 0
for t in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4000):
    n = random.randint(1, 120); m = random.randint(1, n); R = random.randint(1, 14)
    if t % 5 == 0: m = n
    cells, Rp, P, S = cellset(n, m, R)
    rf = rows_formula(n, m, Rp)
    for i in range(n):
        js = sorted(j for (ii, j) in cells if ii == i)
        ok = js and js == list(range(js[0], js[-1] + 1)) and (js[0], js[-1]) == rf[i]
        if not ok:
            bad += 1
            if bad < 10: print("MISMATCH n,m,R,Rp", n, m, R, Rp, "row", i, "set", js[:1], js[-1:], "formula", rf[i], "contig", js == list(range(js[0], js[-1]+1)) if js else None)
            break
print("bad", bad)
