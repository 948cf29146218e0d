"""CPU model of the fast bin lookup of csrc/track_common.cuh (bin_guess / hist_axis_setup): the
same fp32 operations restated in numpy, checked against a search of the edge table
(torch.histogramdd's rule: left-closed bins, the last one right-closed).

The kernel trusts floor(t), t = fma(x, inv, off), wherever frac(t) is farther than
    margin = 1.25 d + 1.001 eps + 2^-20,   eps = n a + b + (n + 1) 2^-24 <= n 2^-22
from 0 and 1 (d: measured deviation of the table from the ideal grid through its end points, in
bins; a, b: the measured errors of inv and off for the table at hand) and lets the table decide
everywhere else.  The derivation is in the header;
this test throws tables at it for which the bound is tight -- one end of the screen at zero, so
that |e0| or |en| equals the extent and u = n at the far end -- and coordinates that sit a few
ulp to a few margins from every edge."""
import numpy as np
import pytest

F32 = np.float32


def _setup(e):
    """hist_axis_setup_staged + hist_axis_finish: e0, en, inv, margin, ok, eps from the fp32 table
    e (n + 1 entries).  eps = n a + b + (n + 1) 2^-24 with a, b the measured errors of inv and
    off for THIS table."""
    n = len(e) - 1
    f0, fn = e[0], e[-1]
    step = (np.float64(fn) - np.float64(f0)) / n
    ideal = np.float64(f0) + step * np.arange(n + 1, dtype=np.float64)
    d = F32(np.max(np.abs(e.astype(np.float64) - ideal)) / step)
    inv = F32(F32(n) / F32(fn - f0))
    off = F32(-f0 * inv)
    inv_true = np.float64(n) / (np.float64(fn) - np.float64(f0))
    a = abs(np.float64(inv) - inv_true) / inv_true
    b = abs(np.float64(off) + np.float64(f0) * np.float64(inv))
    eps = F32(n * a + b + (n + 1) * 2.0 ** -24)
    m = F32(F32(1.25) * d + F32(1.001) * eps + F32(2.0 ** -20))
    ok = d < 0.2 and m < 0.25 and step > 0 and max(abs(f0), abs(fn)) <= F32(fn - f0)
    return f0, fn, inv, m, ok, eps


def _fast_bins(x, f0, fn, inv, m):
    """bin_guess: returns (bin or -1, near).  fma emulated in fp64 (the product of two fp32
    numbers is exact there)."""
    off = F32(-f0 * inv)
    t = (x.astype(np.float64) * np.float64(inv) + np.float64(off)).astype(F32)
    i = np.floor(t).astype(np.int64)
    fr = (t - i.astype(F32)).astype(F32)
    valid = (x >= f0) & (x <= fn)
    near = valid & ~(np.abs((fr - F32(0.5)).astype(F32)) < F32(F32(0.5) - m))
    return np.where(valid, i, -1), near


def _table_bins(x, e):
    n = len(e) - 1
    j = np.searchsorted(e, x, side="right") - 1
    j = np.where(x == e[-1], n - 1, j)
    return np.where((x >= e[0]) & (x <= e[-1]), j, -1)


def _candidates(e, rng):
    e64 = e.astype(np.float64)
    n = len(e) - 1
    step = (e64[-1] - e64[0]) / n
    m = n * 2.0 ** -22 + 2.0 ** -20
    pts = [e]
    up, dn = e.copy(), e.copy()
    for _ in range(12):
        up, dn = np.nextafter(up, F32(np.inf)), np.nextafter(dn, F32(-np.inf))
        pts += [up.copy(), dn.copy()]
    for f in (0.25, 0.5, 0.75, 0.9, 0.99, 1.0, 1.01, 1.1, 1.25, 1.5, 2.0, 3.0):
        pts += [(e64 + f * m * step).astype(F32), (e64 - f * m * step).astype(F32)]
    pts.append(rng.uniform(e64[0], e64[-1], 200_000).astype(F32))
    return np.concatenate(pts)


TABLES = [
    # (n, lo, hi): the reference's screens are symmetric; one-sided ones make |e0| or |en| equal
    # to the extent, the case the error bound is derived for
    (2448, -4.3437e-3, 4.3437e-3), (2040, -2.5503e-3, 2.5503e-3),
    (2448, 0.0, 8.6874e-3), (2448, -8.6874e-3, 0.0), (2040, 1.0e-3, 2.0e-3),
    (32767, 0.0, 1.0), (32767, -1.0, 0.0), (32767, -0.37, 0.63), (4096, 0.0, 3.0), (1000, -7.0, 0.0),
    (17, 0.0, 1e-6), (1, -1.0, 1.0), (3, 0.0, 3e-3),
]


@pytest.mark.parametrize("n,lo,hi", TABLES)
def test_fast_path_never_disagrees_with_the_table(n, lo, hi):
    rng = np.random.default_rng(n)
    e = np.linspace(np.float64(lo), np.float64(hi), n + 1).astype(F32)  # torch.linspace-like
    f0, fn, inv, m, ok, _ = _setup(e)
    if not ok:
        pytest.skip("table takes the searched path in the kernel")
    x = _candidates(e, rng)
    fast, near = _fast_bins(x, f0, fn, inv, m)
    want = _table_bins(x, e)
    trusted = ~near
    bad = trusted & (fast != want)
    assert not bad.any(), (f"{int(bad.sum())} trusted lookups wrong, first x = {x[bad][0]!r}: "
                           f"fast {fast[bad][0]}, table {want[bad][0]}")
    # the table decides only a small share of an interior beam's lookups
    interior = x[-200_000:]
    _, near_i = _fast_bins(interior, f0, fn, inv, m)
    assert near_i.mean() <= 2.5 * float(m) + 2e-4


def test_the_candidates_reach_the_error_zone():
    """Negative control: floor(t) alone (margin 0) gets lookups next to the edges wrong on every
    kind of table -- the candidates do sit where the rounding of t decides."""
    for n, lo, hi in [(32767, -0.37, 0.63), (2448, -4.3437e-3, 4.3437e-3), (32767, 0.0, 1.0),
                      (30011, -0.911, 0.0893), (20000, -0.0131, 0.9771)]:
        e = np.linspace(np.float64(lo), np.float64(hi), n + 1).astype(F32)
        f0, fn, inv, _, ok, _ = _setup(e)
        assert ok
        x = _candidates(e, np.random.default_rng(1))
        fast, _ = _fast_bins(x, f0, fn, inv, F32(0.0))
        assert (fast != _table_bins(x, e)).any()


@pytest.mark.parametrize("n,lo,hi", TABLES)
def test_error_of_t_stays_below_the_bound(n, lo, hi):
    """|t - u| <= eps = n a + b + (n + 1) 2^-24 <= n 2^-22 for every coordinate inside the table
    (u: the exact position on the ideal grid; a, b: this table's errors of inv and off) -- the
    inequality the margin is built on, checked directly."""
    e = np.linspace(np.float64(lo), np.float64(hi), n + 1).astype(F32)
    f0, fn, inv, _, ok, eps = _setup(e)
    if not ok:
        pytest.skip("table takes the searched path in the kernel")
    x = _candidates(e, np.random.default_rng(2))
    x = x[(x >= f0) & (x <= fn)]
    off = F32(-f0 * inv)
    t = (x.astype(np.float64) * np.float64(inv) + np.float64(off)).astype(F32)
    step = (np.float64(fn) - np.float64(f0)) / n
    u = (x.astype(np.float64) - np.float64(f0)) / step
    err = np.max(np.abs(t.astype(np.float64) - u))
    assert err <= float(eps) <= 1.001 * n * 2.0 ** -22 + 2.0 ** -23
