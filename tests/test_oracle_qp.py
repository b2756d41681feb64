"""Pins the oracle's stage 3 (CPU only): the assembled H = D^T D against the closed form and
against a dense numpy D^T D; doStep/runQP against an independent numpy fp32 restatement."""
import numpy as np

from conftest import random_blobs


def band_structures(O, v, bs=4.0):
    m = O.dfield(v)
    band, tight = O.create_bands(m, bs)
    idmap = -np.ones(v.shape, np.int64)
    pos = np.argwhere(band != 0)          # argwhere is (x,y,z)-lexicographic == reference order
    idmap[tuple(pos.T)] = np.arange(len(pos))
    return m, band, tight, pos, idmap


def dense_H(v_shape, tight, pos, idmap):
    n = len(pos)
    rows = []
    for t in np.argwhere(tight != 0):
        for a in range(3):
            e = np.zeros(3, int); e[a] = 1
            r = np.zeros(n)
            r[idmap[tuple(t)]] = -2
            r[idmap[tuple(t - e)]] = 1
            r[idmap[tuple(t + e)]] = 1
            rows.append(r)
    D = np.array(rows)
    return D.T @ D


def csc_to_dense(q):
    n = q["n"]
    R = np.zeros((n, n))
    for c in range(n):
        for k in range(q["jc"][c], q["jc"][c + 1]):
            R[q["ir"][k], c] = q["pr"][k]
    return R


def test_H_equals_dense_DtD_small(O, synth):
    v = synth.sphere(20, 3.0)
    m, band, tight, pos, idmap = band_structures(O, v)
    q = O.prime_qp(v, m, band, tight)
    H = dense_H(v.shape, tight, pos, idmap)
    R = csc_to_dense(q)
    assert np.array_equal(R, H - np.diag(np.diag(H)))
    dg = np.diag(H)
    assert np.array_equal(q["invdg"], (1.0 / dg.astype(np.float32).astype(np.float64)).astype(np.float32))
    # rows ascending inside each column (sorted CSC contract, quadprog.cpp:11-40)
    for c in range(q["n"]):
        col = q["ir"][q["jc"][c]:q["jc"][c + 1]]
        assert np.all(np.diff(col) > 0)


def closed_form(tight, pos, idmap):
    """d(v), c(v,v+-e), c(v,v+-2e) from the tight mask."""
    T = np.pad(tight != 0, 2).astype(np.int64)
    P = pos + 2
    t0 = T[tuple(P.T)]
    out = {}
    diag = np.zeros(len(pos), np.int64)
    for a in range(3):
        e = np.zeros(3, int); e[a] = 1
        tm = T[tuple((P - e).T)]
        tp = T[tuple((P + e).T)]
        diag += 4 * t0 + tm + tp
        out[(a, -1)] = -2 * (t0 + tm)
        out[(a, +1)] = -2 * (t0 + tp)
        out[(a, -2)] = tm
        out[(a, +2)] = tp
    return diag, out


def test_closed_form_matches_assembled(O, synth):
    for v in (synth.sphere(40, 11.0), random_blobs((36, 38, 40), 4), synth.box(40, (9, 7, 5), None, 25.0, 10.0)):
        m, band, tight, pos, idmap = band_structures(O, v)
        q = O.prime_qp(v, m, band, tight)
        diag, coef = closed_form(tight, pos, idmap)
        assert diag.min() >= 1 and diag.max() <= 18
        assert np.array_equal(q["invdg"], (1.0 / diag.astype(np.float64)).astype(np.float32))
        # rebuild the CSC from the closed form in ascending neighbour-id order
        order = [(0, -2), (0, -1), (1, -2), (1, -1), (2, -2), (2, -1), (2, 1), (2, 2), (1, 1), (1, 2), (0, 1), (0, 2)]
        padded = np.pad(idmap, 2, constant_values=-1)
        P = pos + 2
        ir, pr, jc = [], [], [0]
        cols = []
        for (a, s) in order:
            e = np.zeros(3, int); e[a] = s
            cols.append((padded[tuple((P + e).T)], coef[(a, s)]))
        for i in range(len(pos)):
            for ids, c in cols:
                if c[i] != 0:
                    assert ids[i] >= 0           # non-zero coefficient => neighbour is in the band
                    ir.append(ids[i]); pr.append(c[i])
            jc.append(len(ir))
        assert np.array_equal(q["jc"], np.array(jc, np.int32))
        assert np.array_equal(q["ir"], np.array(ir, np.int32))
        assert np.array_equal(q["pr"], np.array(pr, np.float32))


def np_run_qp(q, iters):
    """Independent fp32 restatement of doStep/runQP (quadprog.cpp:45-91), row by row."""
    f32 = np.float32
    x = q["x0"].copy()
    jc, ir, pr = q["jc"], q["ir"], q["pr"]
    for _ in range(iters):
        new = np.empty_like(x)
        for r in range(q["n"]):
            res = f32(0)
            for k in range(jc[r], jc[r + 1]):
                res = f32(res + f32(pr[k] * x[ir[k]]))
            t = f32(f32(x[r] - f32(res * q["invdg"][r])) * f32(0.5))
            if t < q["lb"][r]:
                t = q["lb"][r]
            if t > q["ub"][r]:
                t = q["ub"][r]
            new[r] = t
        x = new
    return x


def test_run_qp_against_numpy(O, synth):
    v = synth.sphere(24, 5.0)
    m, band, tight, pos, idmap = band_structures(O, v)
    q = O.prime_qp(v, m, band, tight)
    for iters in (1, 2, 5):
        assert np.array_equal(O.run_qp(q, iters=iters), np_run_qp(q, iters))


def test_bounds_and_x0(O, synth):
    v = synth.sphere(40, 11.0)
    m, band, tight, pos, idmap = band_structures(O, v)
    q = O.prime_qp(v, m, band, tight)
    inside = v[tuple(pos.T)] == 1
    mg = m[tuple(pos.T)]
    assert np.array_equal(q["lb"], np.where(inside, mg, -1000).astype(np.float32))
    assert np.array_equal(q["ub"], np.where(inside, 1000, -mg).astype(np.float32))
    assert np.array_equal(q["x0"], np.where(inside, mg, -mg).astype(np.float32))
    x = O.run_qp(q, iters=500)
    assert np.all(x >= q["lb"]) and np.all(x <= q["ub"])
    assert np.all(np.isfinite(x))


def test_feature_reset_semantics(O, synth):
    v = synth.sphere(24, 5.0)
    m, band, tight, pos, idmap = band_structures(O, v)
    q = O.prime_qp(v, m, band, tight)
    base = O.run_qp(q, iters=3)
    got = O.run_qp(q, iters=3, feature_idx=np.array([2, 9, 40], np.int32), using_features=True)
    exp = base.copy(); exp[[2, 9, 40]] = 0.2
    assert np.array_equal(got, exp)
    # a -1 (feature voxel outside the band) sorts first and blocks every reset (quadprog.cpp:94-100)
    got = O.run_qp(q, iters=3, feature_idx=np.array([-1, 2, 9], np.int32), using_features=True)
    assert np.array_equal(got, base)


def test_optimize_assembles_F(O, synth):
    v = synth.sphere(40, 11.0)
    F, info = O.optimize(v, return_info=True)
    m, band, tight, pos, idmap = band_structures(O, v)
    q = O.prime_qp(v, m, band, tight)
    x = O.run_qp(q, iters=500)
    expect = np.where(v == 1, 5.0, -5.0).astype(np.float32)
    expect[tuple(pos.T)] = x
    assert np.array_equal(F, expect)
    assert info["n_band"] == len(pos)
