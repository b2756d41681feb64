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


@pytest.mark.gpu
def test_batch_stacked_equals_one_by_one(ctx, synth):
    """mrb_reconstruct_batch runs a batch as ONE stacked grid when the objects' x-borders are empty; every object must
    come out exactly as from mrb_reconstruct on that object alone (odd x size: bricks straddle the object seams)."""
    vols = np.stack([synth.random_object(2000 + i, n=47) for i in range(9)])
    meshes = ctx.reconstruct_batch(vols)
    for i, (verts, faces) in enumerate(meshes):
        v1, f1 = ctx.reconstruct(vols[i])
        assert np.array_equal(faces, f1) and np.array_equal(verts, v1), i
    # an object whose shape comes too close to its x-border cannot be stacked; alone it is the reference's "band touches the
    # border" exit, and the batch must report it the same way, naming the object
    vols2 = vols.copy()
    vols2[4] = 0.0
    vols2[4, 4:20, 15:30, 15:30] = 1.0
    with pytest.raises(Exception, match="object 4.*border"):
        ctx.reconstruct_batch(vols2)
    # an empty object in the middle of a stack, and iso != 0
    vols3 = vols.copy()
    vols3[2] = 0.0
    meshes3 = ctx.reconstruct_batch(vols3, isolevel=0.25)
    for i, (verts, faces) in enumerate(meshes3):
        v1, f1 = ctx.reconstruct(vols3[i], isolevel=0.25)
        assert np.array_equal(faces, f1) and np.array_equal(verts, v1), i
    assert len(meshes3[2][1]) == 0
