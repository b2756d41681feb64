"""The torch generators used by bench.py's slab record produce exactly the numpy volumes."""
import numpy as np


def test_torch_generators_equal_numpy(synth):
    for n, r, c in ((52, 17.0, None), (40, 9.3, (17.2, 20.9, 14.4)), ((30, 26, 33), 7.77, (14.5, 12.0, 16.0))):
        assert np.array_equal(synth.sphere(n, r, c), synth.sphere_t(n, r, c).numpy())
    assert np.array_equal(synth.config_volume(1), synth.config_volume_t(1).numpy())
    rng = np.random.default_rng(4)
    for _ in range(8):
        r = rng.uniform(5, 10)
        c = [rng.uniform(20, 40) for _ in range(3)]
        assert np.array_equal(synth.sphere(64, r, c), synth.sphere_t(64, r, c).numpy())
