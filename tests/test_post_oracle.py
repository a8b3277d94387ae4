"""CPU checks of the post-processing restatement (oracle/post_oracle.py) on synthetic films:
properties any correct implementation of the stated definition has."""
import numpy as np

from fs2_b200 import abi
from oracle import post_oracle


def _film(h=24, w=40, seed=0):
    rng = np.random.default_rng(seed)
    mean = rng.random((h, w, 3), dtype=np.float32)
    normal = np.zeros((h, w, 3), np.float32)
    normal[..., 1] = 1.0
    depth = np.full((h, w), 50.0, np.float32)
    ids = np.ones((h, w), np.uint64)
    ids[:, w // 2:] = 2  # two materials side by side
    return mean, normal, depth, ids


def test_constant_image_is_a_fixed_point():
    mean, normal, depth, ids = _film()
    mean[:] = (0.25, 0.5, 0.75)
    rgb, rgb8 = post_oracle.post(mean, normal, depth, ids, abi.post_params(denoise_passes=4))
    assert np.allclose(rgb, mean, rtol=1e-6)
    assert (rgb8 == rgb8[0, 0]).all()


def test_denoise_never_mixes_material_ids():
    mean, normal, depth, ids = _film()
    mean[:, : mean.shape[1] // 2] = (1.0, 0.0, 0.0)
    mean[:, mean.shape[1] // 2:] = (0.0, 0.0, 1.0)
    rgb, _ = post_oracle.post(mean, normal, depth, ids, abi.post_params(denoise_passes=3))
    assert np.array_equal(rgb, mean)


def test_denoise_reduces_noise_and_keeps_the_mean():
    mean, normal, depth, ids = _film(48, 64)
    mean = (0.5 + 0.1 * (mean - 0.5)).astype(np.float32)
    rgb, _ = post_oracle.post(mean, normal, depth, ids, abi.post_params(denoise_passes=3))
    assert rgb.std() < 0.5 * mean.std()
    assert abs(rgb.mean() - mean.mean()) < 0.01


def test_stages_off_is_identity_and_bloom_adds_light():
    mean, normal, depth, ids = _film()
    rgb, _ = post_oracle.post(mean, normal, depth, ids, abi.post_params(denoise_passes=0))
    assert np.array_equal(rgb, mean)
    mean[:] = 0.1
    mean[10, 20] = 8.0  # one emitter pixel
    p = abi.post_params(denoise_passes=0, bloom_threshold=1.0, bloom_strength=0.5, bloom_radius=4,
                        flare_strength=0.5, flare_ghosts=2)
    rgb, _ = post_oracle.post(mean, normal, depth, ids, p)
    assert (rgb >= mean - 1e-7).all()
    assert rgb[10, 23, 0] > mean[10, 23, 0]  # halo next to the emitter
    h, w = depth.shape
    assert rgb[h - 1 - 10, w - 1 - 20, 0] > 0.1 + 1e-4  # full-scale ghost mirrored through the centre
    assert rgb[0, 0, 0] == np.float32(0.1)  # far from both
