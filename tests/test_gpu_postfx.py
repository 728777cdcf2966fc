"""GPU: canvas compositing and post-FX kernels (vt_fx_compose / vt_fx_draw) vs the numpy restatement of
Window.draw_tile / Window.draw (oracle/postfx_np.py), bit for bit.  pygame's own rounding is unpinned."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cam(w, h):
    from python_raytracer_b200 import GpuCamera, Settings
    return GpuCamera(Settings.from_object(dict(width=w, height=h, chunk_size=16)))


def _rand_rgba(rng, h, w, zero_alpha=0.2):
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    img[rng.random((h, w)) < zero_alpha, 3] = 0
    img[rng.random((h, w)) < 0.1, 3] = 255
    return img


def test_draw_tile_compositing_matches_restatement():
    from oracle import postfx_np as ref
    from python_raytracer_b200.postfx import Canvas
    rng = np.random.default_rng(3)
    cam = _cam(96, 64)
    cv = Canvas(cam)
    expect = np.zeros((64, 96, 4), np.uint8)
    for k in range(5):                      # fresh canvas (alpha 0), then repeated blits = motion blur accumulation
        tile = _rand_rgba(rng, 64, 96)
        cv.draw_tile(torch.from_numpy(tile).cuda())
        expect = ref.compose(expect, tile)
        assert np.array_equal(cv.canvas.cpu().numpy(), expect)
    cam.close()


@pytest.mark.parametrize("fx", [
    dict(),                                                                                     # nothing on, scale only
    dict(spill=0.1),
    dict(iris=0.5, iris_time=0.5, iris_state=0.4),
    dict(iris=0.5, iris_time=0.5, iris_state=-0.7),
    dict(bloom=0.5, bloom_blur=4),
    dict(spill=0.1, iris=0.5, iris_time=0.5, iris_state=0.25, bloom=0.5, bloom_blur=4, subsamples=0.5, smooth=0.25),   # shipped config.cfg
    dict(spill=0.3, bloom=0.25, bloom_blur=3, subsamples=1.0, smooth=1.0),
    dict(smooth=0.5, subsamples=0.25),
])
@pytest.mark.parametrize("scale", [1, 3])
def test_draw_matches_restatement(fx, scale):
    from oracle import postfx_np as ref
    from python_raytracer_b200.postfx import Canvas
    rng = np.random.default_rng(11)
    w, h = 64, 48
    cam = _cam(w, h)
    cv = Canvas(cam, scale=scale)
    img = _rand_rgba(rng, h, w, zero_alpha=0.05)
    cv.canvas.copy_(torch.from_numpy(img))
    fx = dict(fx)
    cv.iris = fx.pop("iris_state", 0.0)
    out = cv.draw(**fx).cpu().numpy()
    exp, color = ref.draw(img, w, h, w * scale, h * scale, spill=fx.get("spill", 0.0),
                          iris_on=bool(fx.get("iris") and fx.get("iris_time")), iris=cv.iris, bloom=fx.get("bloom", 0.0),
                          bloom_blur=fx.get("bloom_blur", 0.0), subsamples=fx.get("subsamples", 0.0), smooth=fx.get("smooth", 0.0))
    assert cv.average == color
    assert out.shape == exp.shape
    assert np.array_equal(out, exp), "max diff %d" % int(np.abs(out.astype(int) - exp.astype(int)).max())
    if fx.get("iris") and fx.get("iris_time"):
        assert cv.iris_target == 1 - (max(color[:3]) / 255) * 2
    cam.close()


def test_rendered_frame_through_compositing_and_fx():
    """End to end on the device: shipped 64x48 frame -> tile() -> draw_tile -> draw with the shipped config.cfg effects."""
    from oracle import postfx_np as ref
    from python_raytracer_b200 import GpuCamera, Settings
    from python_raytracer_b200.postfx import Canvas
    from tests.golden_util import load_case
    c = load_case("c1_shipped")
    sd = dict(c["settings"]); sd["threads"] = 1

    def bg(ray, settings):
        raise AssertionError

    cam = GpuCamera(Settings.from_object(sd), background_function=bg, background=bg)
    cam.set_world_dense(c["origin"], c["cells"], c["chunk_res"], c["materials"])
    cam.pos, cam.rot, cam.lens = c["cam_pos"], c["cam_rot"], float(c["cam_lens"])
    cv = Canvas(cam, scale=16)
    frame = cam.render(want=("rgba8",))["rgba8"]
    cv.draw_tile(frame); cv.draw_tile(frame)
    exp_canvas = ref.compose(ref.compose(np.zeros((48, 64, 4), np.uint8), frame.cpu().numpy()), frame.cpu().numpy())
    assert np.array_equal(cv.canvas.cpu().numpy(), exp_canvas)
    out = cv.draw(spill=0.1, iris=0.5, iris_time=0.5, bloom=0.5, bloom_blur=4, subsamples=0.5, smooth=0.25).cpu().numpy()
    exp, _ = ref.draw(exp_canvas, 64, 48, 1024, 768, spill=0.1, iris_on=True, iris=0.0, bloom=0.5, bloom_blur=4, subsamples=0.5, smooth=0.25)
    assert out.shape == (768, 1024, 4) and np.array_equal(out, exp)
    cam.close()
