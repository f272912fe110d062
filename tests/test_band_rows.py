"""CPU: the per-row closed form of the banded cell set that the CUDA row-sweep kernels use
(rawalign_b200/csrc/dtw_core.cuh) equals the anti-diagonal enumeration of the reference
(dtw.cpp:298-302, :352-366, :416-420), and its size equals ra_dtw_cells / the oracle."""
import random


def cellset(n, m, R):
    extra = ((n - m) * R + n - 1) // n
    Rp = R + extra
    P = Rp + (1 if Rp % 2 == 0 else 0); S = Rp + (1 if Rp % 2 == 1 else 0)
    cells = {(0, 0)}

    def seg(si, sj, ln):
        for o in range(ln):
            i, j = si - o, sj + o
            if 0 <= i < n and 0 <= j < m:
                cells.add((i, j))
    r = 0
    for c in range(1, n):
        if (r + 1) * n <= m * c:
            r += 1; seg(c + S // 2 - 1, r - S // 2, S)
        seg(c + P // 2, r - P // 2, P)
    return cells, Rp


def rows_formula(n, m, Rp):
    odd = Rp % 2; h = (Rp - 1) // 2 if odd else Rp // 2
    out = []
    for i in range(n):
        ih = i + h + odd; il = i - h - odd
        jhi = m - 1 if ih > n - 1 else min(m - 1, (m * ih) // n + h)
        jlo = 0 if il < 0 else max(0, (m * il) // n - h)
        out.append((jlo, jhi))
    return out


def test_row_form_equals_antidiagonal_enumeration():
    from oracle import Oracle
    orc = Oracle()
    rnd = random.Random(7)
    for t in range(700):
        n = rnd.randint(1, 90); m = rnd.randint(1, n); R = rnd.randint(1, 12)
        if t % 5 == 0:
            m = n
        cells, Rp = cellset(n, m, R)
        rf = rows_formula(n, m, Rp)
        tot = 0
        for i in range(n):
            js = sorted(j for (ii, j) in cells if ii == i)
            assert js == list(range(js[0], js[-1] + 1)), (n, m, R, i)
            assert (js[0], js[-1]) == rf[i], (n, m, R, i)
            if i:
                assert rf[i][0] >= rf[i - 1][0] and rf[i][1] >= rf[i - 1][1] and rf[i][0] - rf[i - 1][0] <= 1
            tot += len(js)
        assert tot == len(cells) == orc.dtw_cells(n, m, R)
        assert rf[-1][1] == m - 1 and rf[0][0] == 0
