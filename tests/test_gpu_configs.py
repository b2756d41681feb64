"""GPU parity at the BASELINE configs: direct comparison with the oracle where it finishes in
seconds (configs 1, 2, 3, 5), size-independent properties at 512^3 (config 4)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def euler_and_manifold(verts, faces):
    e = np.sort(np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]]), axis=1)
    uniq, cnt = np.unique(e, axis=0, return_counts=True)
    return len(verts) - len(uniq) + len(faces), bool(np.all(cnt == 2))


def test_config2_172_box_union_sphere(ctx, O, synth):
    v = synth.config_volume(2)
    F_ref, info = O.optimize(v, brute_edt=False, return_info=True)
    verts_ref, faces_ref = O.mcubes(F_ref, 0.0)
    F, st = ctx.optimize(v, return_stats=True)
    assert st["n_band"] == info["n_band"]
    assert np.array_equal(F, F_ref)
    verts, faces = ctx.reconstruct(v)
    assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref)


@pytest.mark.parametrize("shape", ["sphere", "torus"])
def test_config3_256(ctx, O, synth, shape):
    v = synth.sphere(256, 110.0) if shape == "sphere" else synth.torus(256, 80.0, 30.0)
    F_ref, info = O.optimize(v, brute_edt=False, return_info=True)
    verts_ref, faces_ref = O.mcubes(F_ref, 0.0)
    verts, faces, st = ctx.reconstruct(v, return_stats=True)
    assert st["n_band"] == info["n_band"]
    if shape == "sphere":
        assert st["n_band"] == 1587528 and st["n_tight"] == 1334320
    assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref)
    chi, manifold = euler_and_manifold(verts, faces)
    assert manifold and chi == (2 if shape == "sphere" else 0)


def test_config3_wide_band_8(ctx, O, synth):
    v = synth.sphere(256, 110.0)
    F_ref, info = O.optimize(v, band_size=8, iters=60, brute_edt=False, return_info=True)
    F, st = ctx.optimize(v, band_size=8, iters=60, return_stats=True)
    assert st["n_band"] == info["n_band"] == 2833384                      # "wide" band of SURVEY.md section 8
    assert np.array_equal(F, F_ref)


def test_wide_band_beyond_byte_window(ctx, O, synth):
    # band_size > 15 leaves the byte-grid path: exact float EDT + float morphology + z-strip kernel
    v = synth.sphere(96, 24.0)
    F_ref = O.optimize(v, band_size=17, iters=25, brute_edt=False)
    F = ctx.optimize(v, band_size=17, iters=25)
    assert np.array_equal(F, F_ref)


def test_config4_512_properties(ctx, O, synth):
    v = synth.config_volume(4)
    verts, faces, st = ctx.reconstruct(v, return_stats=True)
    # band membership against the oracle's stage 1+2 (exact, seconds), the rest through invariants
    m = O.dfield(v, brute=False)
    band, tight = O.create_bands(m, 4.0)
    assert st["n_band"] == int(band.sum()) and st["n_tight"] == int(tight.sum())
    F = ctx.optimize(v)
    inb, inside = band != 0, v == 1
    assert np.all(F[inb & inside] >= m[inb & inside]) and np.all(F[inb & ~inside] <= -m[inb & ~inside])
    assert np.all(F[~inb & inside] == 5.0) and np.all(F[~inb & ~inside] == -5.0)
    chi, manifold = euler_and_manifold(verts, faces)
    assert manifold and chi == 2                                           # sphere with bumps: one closed surface
    flat = faces.reshape(-1)
    first = np.full(len(verts), len(flat), np.int64)
    np.minimum.at(first, flat, np.arange(len(flat)))
    assert np.all(np.diff(first) > 0)                                      # vertex ids in first-appearance order
    # MC of the GPU embedding == oracle MC of the same embedding (stage 4 alone is cheap on the CPU)
    verts_ref, faces_ref = O.mcubes(F, 0.0)
    assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref)


def test_config5_batch_of_objects(ctx, O, synth):
    vols = np.stack([synth.random_object(1000 + i) for i in range(6)])
    meshes = ctx.reconstruct_batch(vols)
    assert len(meshes) == 6
    for i, (verts, faces) in enumerate(meshes):
        verts_ref, faces_ref = O.mcubes(O.optimize(vols[i]), 0.0)
        assert np.array_equal(faces, faces_ref) and np.array_equal(verts, verts_ref), i
