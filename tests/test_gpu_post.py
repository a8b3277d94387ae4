"""film post-processing on the device (fs_film_post: guided a-trous
denoise, bloom, flare) against its CPU restatement oracle/post_oracle.py.

The reference has no code for these stages (only the guide buffers, source/frame.h:86-91), so
there is no upstream parity target; the checks are
  * device == numpy restatement within a stated float tolerance (expf / powf differ by a few
    ulp between libdevice and numpy; sums run in the same order),
  * with every stage off the output is the film's own mean / rgb8, bit for bit,
  * the film is not modified,
  * denoising a low-spp image moves it towards the converged image,
  * argument errors are reported, not clamped."""
import numpy as np
import pytest

from fs2_b200 import abi
from oracle import post_oracle
from refdata import golden, rel_rmse

pytestmark = pytest.mark.gpu

RTOL, ATOL = 2e-4, 2e-6  # float32 image values, exp / pow implementation differences


def _render(ctx, name, w, h, spp, seed=3):
    g = golden(name)
    ctx.upload(g.scene)
    ctx.resize(w, h)
    ctx.trace(g.cam, spp, seed=seed)
    return ctx.read_film(aux=True)


def _oracle(film, w, h, p):
    return post_oracle.post(film["mean"].reshape(h, w, 3), film["normal"].reshape(h, w, 3),
                            film["depth"].reshape(h, w), film["mat_id"].reshape(h, w), p)


@pytest.mark.parametrize("name,w,h", [("demo", 96, 64), ("mesh_tiny", 80, 48), ("glass_small", 67, 45)])
@pytest.mark.parametrize("stages", ["denoise", "glow", "all"])
def test_post_matches_cpu_restatement(ctx, name, w, h, stages):
    film = _render(ctx, name, w, h, 8)
    p = abi.post_params(denoise_passes=3 if stages != "glow" else 0,
                        bloom_threshold=0.6,
                        bloom_strength=0.5 if stages != "denoise" else 0.0, bloom_radius=5,
                        flare_strength=0.25 if stages != "denoise" else 0.0,
                        flare_ghosts=3 if stages != "denoise" else 0)
    got = ctx.post_film(p)
    ref_rgb, ref_rgb8 = _oracle(film, w, h, p)
    assert np.isfinite(got["rgb"]).all()
    assert np.allclose(got["rgb"].reshape(h, w, 3), ref_rgb, rtol=RTOL, atol=ATOL)
    # display bytes: the quantiser may round a value within tolerance to the neighbouring code
    d = np.abs(got["rgb8"].reshape(h, w, 3).astype(int) - ref_rgb8.astype(int))
    assert d.max() <= 1 and (d != 0).mean() < 0.01


def test_post_with_all_stages_off_is_the_film(ctx):
    film = _render(ctx, "demo", 64, 40, 4)
    got = ctx.post_film(abi.post_params(denoise_passes=0))
    assert got["rgb"].tobytes() == film["mean"].tobytes()
    assert got["rgb8"].tobytes() == film["rgb8"].tobytes()
    again = ctx.read_film(aux=True)
    for k in ("mean", "count", "rgb8", "normal", "depth", "mat_id"):
        assert again[k].tobytes() == film[k].tobytes(), k  # film untouched


def test_denoise_moves_low_spp_towards_converged(ctx):
    w, h = 160, 96
    ref = _render(ctx, "demo", w, h, 1024, seed=11)["mean"]
    ctx.reset()
    noisy = _render(ctx, "demo", w, h, 4, seed=5)["mean"]
    den = ctx.post_film(abi.post_params(denoise_passes=4))["rgb"]
    e0, e1 = rel_rmse(noisy, ref), rel_rmse(den, ref)
    print("4 spp vs 1024 spp: rel RMSE %.4f -> %.4f after 4 a-trous passes" % (e0, e1))
    assert e1 < 0.7 * e0
    # energy is preserved to a few percent (normalised weights)
    assert abs(den.mean() - noisy.mean()) < 0.05 * noisy.mean()


def test_post_rejects_bad_arguments(ctx):
    _render(ctx, "demo", 32, 20, 1)
    for bad in (abi.post_params(denoise_passes=9), abi.post_params(denoise_passes=-1),
                abi.post_params(bloom_strength=1.0, bloom_radius=0),
                abi.post_params(flare_strength=1.0, flare_ghosts=17),
                abi.post_params(denoise_passes=1, sigma_depth=0.0)):
        with pytest.raises(RuntimeError):
            ctx.post_film(bad)
    # still usable afterwards
    ctx.post_film(abi.post_params(denoise_passes=1))


def test_post_1080p_timing(ctx):
    """Not a gate on speed: prints the device time of the full chain at the bench resolution."""
    import time
    _render(ctx, "demo", 1920, 1080, 1)
    p = abi.post_params(denoise_passes=5, bloom_strength=0.3, bloom_radius=24,
                        flare_strength=0.1, flare_ghosts=4)
    ctx.post_film(p, rgb=False, rgb8=False)
    t0 = time.perf_counter()
    ctx.post_film(p, rgb=False, rgb8=False)
    dt = time.perf_counter() - t0
    print("fs_film_post 1080p, 5 passes + bloom r=24 + 4 ghosts: %.2f ms" % (dt * 1e3))
    assert dt < 0.5
