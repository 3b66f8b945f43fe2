"""Independent pure-Python restatement of Go 1.16 sort.Sort (src/sort/sort.go) driven through
Less/Swap callbacks, plus Binpack/Spread.Choose on top of it.  Used only to cross-check the C
oracle on small random cases; written against the published algorithm, not against the C code."""


class Cells:
    def __init__(self, free, rl, idx):
        self.a = [[f, r, i] for f, r, i in zip(free, rl, idx)]

    def __len__(self):
        return len(self.a)

    def less(self, i, j):
        return self.a[i][0] + self.a[i][1] * 50 < self.a[j][0] + self.a[j][1] * 50

    def swap(self, i, j):
        self.a[i], self.a[j] = self.a[j], self.a[i]


def _insertion(d, a, b):
    for i in range(a + 1, b):
        j = i
        while j > a and d.less(j, j - 1):
            d.swap(j, j - 1)
            j -= 1


def _sift_down(d, lo, hi, first):
    root = lo
    while True:
        child = 2 * root + 1
        if child >= hi:
            return
        if child + 1 < hi and d.less(first + child, first + child + 1):
            child += 1
        if not d.less(first + root, first + child):
            return
        d.swap(first + root, first + child)
        root = child


def _heap_sort(d, a, b):
    first, lo, hi = a, 0, b - a
    for i in range((hi - 1) // 2, -1, -1):
        _sift_down(d, i, hi, first)
    for i in range(hi - 1, -1, -1):
        d.swap(first, first + i)
        _sift_down(d, lo, i, first)


def _median3(d, m1, m0, m2):
    if d.less(m1, m0):
        d.swap(m1, m0)
    if d.less(m2, m1):
        d.swap(m2, m1)
        if d.less(m1, m0):
            d.swap(m1, m0)


def _do_pivot(d, lo, hi):
    m = (lo + hi) >> 1
    if hi - lo > 40:
        s = (hi - lo) // 8
        _median3(d, lo, lo + s, lo + 2 * s)
        _median3(d, m, m - s, m + s)
        _median3(d, hi - 1, hi - 1 - s, hi - 1 - 2 * s)
    _median3(d, lo, m, hi - 1)
    pivot = lo
    a, c = lo + 1, hi - 1
    while a < c and d.less(a, pivot):
        a += 1
    b = a
    while True:
        while b < c and not d.less(pivot, b):
            b += 1
        while b < c and d.less(pivot, c - 1):
            c -= 1
        if b >= c:
            break
        d.swap(b, c - 1)
        b += 1
        c -= 1
    protect = hi - c < 5
    if not protect and hi - c < (hi - lo) // 4:
        dups = 0
        if not d.less(pivot, hi - 1):
            d.swap(c, hi - 1)
            c += 1
            dups += 1
        if not d.less(b - 1, pivot):
            b -= 1
            dups += 1
        if not d.less(m, pivot):
            d.swap(m, b - 1)
            b -= 1
            dups += 1
        protect = dups > 1
    if protect:
        while True:
            while a < b and not d.less(b - 1, pivot):
                b -= 1
            while a < b and d.less(a, pivot):
                a += 1
            if a >= b:
                break
            d.swap(a, b - 1)
            a += 1
            b -= 1
    d.swap(pivot, b - 1)
    return b - 1, c


def _quick_sort(d, a, b, max_depth):
    while b - a > 12:
        if max_depth == 0:
            _heap_sort(d, a, b)
            return
        max_depth -= 1
        mlo, mhi = _do_pivot(d, a, b)
        if mlo - a < b - mhi:
            _quick_sort(d, a, mlo, max_depth)
            a = mhi
        else:
            _quick_sort(d, mhi, b, max_depth)
            b = mlo
    if b - a > 1:
        for i in range(a + 6, b):
            if d.less(i, i - 6):
                d.swap(i, i - 6)
        _insertion(d, a, b)


def go116_sort(d):
    n = len(d)
    depth, i = 0, n
    while i > 0:
        depth += 1
        i >>= 1
    _quick_sort(d, 0, n, depth * 2)


def go119_sort_small(d):
    _insertion(d, 0, len(d))


def choose(free, rl, demand, spread, sorter=go116_sort):
    """Binpack.Choose / Spread.Choose (rater.go:74-110,127-163) -> (ok, result or partial list)"""
    gp = Cells(list(free), list(rl), range(len(free)))
    dm = Cells(list(demand), [0] * len(demand), range(len(demand)))
    sorter(dm)
    indexes = []
    for j in range(len(dm) - 1, -1, -1):
        if dm.a[j][0] == 0:
            indexes.append(-1)
            continue
        sorter(gp)
        rng = range(len(gp) - 1, -1, -1) if spread else range(len(gp))
        for i in rng:
            if gp.a[i][0] >= dm.a[j][0]:
                indexes.append(gp.a[i][2])
                gp.a[i][0] -= dm.a[j][0]
                break
    if len(indexes) != len(demand):
        return False, indexes
    res = [None] * len(demand)
    for j in range(len(demand) - 1, -1, -1):
        res[dm.a[j][2]] = indexes[len(demand) - 1 - j]
    return True, res
