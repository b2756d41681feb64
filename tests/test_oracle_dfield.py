"""Pins the oracle's stage 1/2 (CPU only): hand-derived known answers, an independent O(N^6)
brute-force EDT in numpy, and brute-force-mode == fast-mode."""
import numpy as np

from conftest import random_blobs


def np_edt_bruteforce(zero_mask):
    """min(100000, min over zero voxels of squared distance) -- independent of the separable algorithm."""
    shape = zero_mask.shape
    zs = np.argwhere(zero_mask)
    out = np.full(shape, 100000, np.int64)
    if len(zs) == 0:
        return out.astype(np.float32)
    g = np.stack(np.meshgrid(*[np.arange(s) for s in shape], indexing="ij"), -1).reshape(-1, 1, 3)
    d2 = ((g - zs[None, :, :]) ** 2).sum(-1).min(1)
    return np.minimum(d2, 100000).reshape(shape).astype(np.float32)


def test_cube_5x5x5_known_answer(O):
    # the only synthetic fixture of the reference: distance_fields/main.cpp:5-23 (5^3 grid, solid 3^3 cube)
    v = np.zeros((5, 5, 5), np.float32)
    v[1:4, 1:4, 1:4] = 1.0
    p = O.fast_perim(v)
    expect = np.zeros((5, 5, 5), np.float32)
    expect[2, 2, 2] = 1.0          # only the cube centre sees a uniform 3x3x3 window
    assert np.array_equal(p, expect)
    sq = O.sqdist(p)
    assert np.array_equal(sq, expect)          # distance 1 at the centre, 0 elsewhere
    assert np.array_equal(O.dfield(v), expect)


def test_single_voxel_known_answer(O):
    v = np.zeros((11, 11, 11), np.float32)
    v[5, 5, 5] = 1.0
    p = O.fast_perim(v)
    # perimeter (value 0) = the 3x3x3 block around the voxel, everything else uniform -> 1
    expect = np.ones_like(v)
    expect[4:7, 4:7, 4:7] = 0.0
    assert np.array_equal(p, expect)
    sq = O.sqdist(p)
    x = np.arange(11)
    d1 = np.maximum(np.abs(x - 5) - 1, 0) ** 2
    expect_sq = d1[:, None, None] + d1[None, :, None] + d1[None, None, :]
    assert np.array_equal(sq, expect_sq.astype(np.float32))
    m = O.dfield(v)
    assert m[5, 5, 0] == 4.0 and m[0, 0, 0] == np.float32(np.sqrt(48.0))


def test_perim_border_clamp(O):
    # uniform grid: no voxel differs from its (clamped) window -> all ones, also on the border
    assert np.all(O.fast_perim(np.ones((4, 5, 6), np.float32)) == 1.0)
    v = np.zeros((4, 5, 6), np.float32)
    v[0, 0, 0] = 1.0
    p = O.fast_perim(v)
    assert p[0, 0, 0] == 0 and p[1, 1, 1] == 0 and p[2, 1, 1] == 1 and p[3, 4, 5] == 1


def test_edt_against_n6_bruteforce(O):
    rng = np.random.default_rng(7)
    for shape, density in (((9, 10, 11), 0.01), ((12, 12, 12), 0.05), ((6, 13, 8), 0.3)):
        g = (rng.random(shape) > density).astype(np.float32)
        ref = np_edt_bruteforce(g == 0)
        assert np.array_equal(O.sqdist(g, brute=True), ref)
        assert np.array_equal(O.sqdist(g, brute=False), ref)
    g = np.ones((5, 5, 5), np.float32)     # no zero voxel -> sentinel everywhere (dfields.cpp:50)
    assert np.all(O.sqdist(g, brute=True) == 100000.0)
    assert np.all(O.sqdist(g, brute=False) == 100000.0)


def test_edt_fast_equals_brute(O, synth):
    for v in (synth.sphere(40, 12.0), random_blobs((30, 34, 38), 2), synth.torus(44, 11.0, 4.0)):
        p = O.fast_perim(v)
        assert np.array_equal(O.sqdist(p, brute=True), O.sqdist(p, brute=False))


def test_edt_clamp_100000(O):
    g = np.ones((340, 2, 2), np.float32)
    g[1, 0, 0] = 0.0
    for brute in (True, False):
        sq = O.sqdist(g, brute=brute)
        assert sq[317, 0, 0] == 316.0 ** 2          # 99856 < 100000
        assert sq[318, 0, 0] == 100000.0            # 317^2 = 100489 -> clamped
        assert sq[339, 1, 1] == 100000.0
        assert sq[1, 1, 1] == 2.0


def test_sqrt_is_correctly_rounded(O):
    a = np.arange(0, 5000, dtype=np.float32).reshape(10, 20, 25)
    assert np.array_equal(O.sqrt(a), np.sqrt(a.astype(np.float64)).astype(np.float32))
    assert np.array_equal(O.sqrt(a), np.sqrt(a))     # fp32 sqrt is correctly rounded too


def test_bands_known_answer_and_properties(O, synth):
    v = synth.sphere(40, 11.0)
    m = O.dfield(v)
    band, tight = O.create_bands(m, 4.0)
    band0 = (m <= 4.0).astype(np.float32)
    assert np.all(band <= band0)                      # opening: band subset of band0
    assert np.all(tight <= band)
    # numpy restatement of erode6 / dilate6 on the interior, border copied (narrowBand.cpp:10-79)
    def erode(g):
        out = g.copy()
        c = g[1:-1, 1:-1, 1:-1]
        r = (c != 0) & (g[:-2, 1:-1, 1:-1] != 0) & (g[2:, 1:-1, 1:-1] != 0) & (g[1:-1, :-2, 1:-1] != 0) & \
            (g[1:-1, 2:, 1:-1] != 0) & (g[1:-1, 1:-1, :-2] != 0) & (g[1:-1, 1:-1, 2:] != 0)
        out[1:-1, 1:-1, 1:-1] = r
        return out
    def dilate(g):
        out = g.copy()
        c = g[1:-1, 1:-1, 1:-1]
        r = (c == 1) | (g[:-2, 1:-1, 1:-1] == 1) | (g[2:, 1:-1, 1:-1] == 1) | (g[1:-1, :-2, 1:-1] == 1) | \
            (g[1:-1, 2:, 1:-1] == 1) | (g[1:-1, 1:-1, :-2] == 1) | (g[1:-1, 1:-1, 2:] == 1)
        out[1:-1, 1:-1, 1:-1] = r
        return out
    assert np.array_equal(tight, erode(band0))
    assert np.array_equal(band, dilate(erode(band0)))
    rng = np.random.default_rng(1)
    g = (rng.random((6, 7, 8)) > 0.3).astype(np.float32)
    assert np.array_equal(O.erode(g), erode(g)) and np.array_equal(O.dilate(g), dilate(g))


def test_find_indexes_order_and_index_map(O):
    g = np.zeros((3, 4, 5), np.float32)
    pts = [(0, 1, 2), (0, 1, 4), (0, 3, 0), (2, 0, 1), (2, 3, 4)]      # already in (x,y,z) lexicographic order
    for p in pts:
        g[p] = 1.0
    idx = O.find_indexes(g)
    # values are the x-fastest linear index z*nx*ny + y*nx + x (grid.cpp:172-174)
    assert idx.tolist() == [z * 3 * 4 + y * 3 + x for (x, y, z) in pts]
    imap = O.index_map(g.shape, idx)
    for i, p in enumerate(pts):
        assert imap[p] == float(i)
    assert (imap == -1).sum() == g.size - len(pts)
    assert len(O.find_indexes(np.zeros((2, 2, 2), np.float32))) == 0


def test_survey_counts_52_sphere(O, synth):
    # 52^3 sphere R=17 -> B = 39000, T = 32488, nnz(R) = 401376
    v = synth.sphere(52, 17.0)
    m = O.dfield(v)
    band, tight = O.create_bands(m, 4.0)
    assert int(band.sum()) == 39000 and int(tight.sum()) == 32488
    q = O.prime_qp(v, m, band, tight)
    assert len(q["ir"]) == 401376
