"""Alias-table importance sampling of the environment map (tinypathtracer_b200/csrc/env.cu, env.cuh) and the optional
next-event estimation built on it (TPT_RENDER_ENV_NEE).

The reference builds a pdf/cdf for its map (src/env_light.cu:10-53) but never samples them (EnvLightSample has no call
site), so there is no reference output to compare with: the checks are the distribution's own invariants
(table mass, sample histogram, pdf consistency, a Monte-Carlo integral) and that NEE leaves the rendered expectation
unchanged while lowering the variance.
"""
import numpy as np
import pytest

from conftest import synthetic_sky

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def built(tpt, ctx, scenes):
    out = {}
    for name in ("ball", "box", "box2"):
        ds = tpt.DeviceScene(ctx, scenes[name])
        bvh = tpt.BVH(ds)
        bvh.construct()
        out[name] = (ds, bvh)
    yield out
    for ds, _ in out.values():
        ds.close()


def spot_sky(w=128, h=64):
    """Dim sky with one small bright patch: the case cosine sampling handles badly."""
    img = np.zeros((h, w, 4), np.uint8)
    img[..., :3] = 6
    img[h // 2:, :, 2] = 12                      # upper hemisphere slightly blue
    img[int(h * 0.70):int(h * 0.76), int(w * 0.10):int(w * 0.13), :3] = 255
    img[..., 3] = 255
    return img


def texel_weights(img):
    h, w = img.shape[:2]
    lum = 0.2126 * img[..., 0].astype(np.float64) + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]
    theta = (1.0 - (np.arange(h) + 0.5) / h) * np.pi
    wgt = lum * np.sin(theta)[:, None]
    return wgt / wgt.sum()


def dir_to_texel(d, w, h):
    u = np.arctan2(d[:, 2], d[:, 0]) / (2 * np.pi)
    u = np.where(u < 0, u + 1, u)
    v = 1.0 - np.arccos(np.clip(d[:, 1], -1, 1)) / np.pi
    x = np.clip(np.floor(u * w).astype(np.int64), 0, w - 1)
    y = np.clip(np.floor(v * h).astype(np.int64), 0, h - 1)
    return x, y


@pytest.mark.parametrize("make", [lambda: synthetic_sky(256, 128), spot_sky])
def test_alias_tables_reproduce_texel_mass(make, tpt, ctx):
    img = make()
    env = tpt.EnvLight(ctx, img)
    prob, alias, mass = env.sampler_tables()
    h, w = img.shape[:2]
    n = w * h
    want = texel_weights(img)
    assert abs(mass.sum(dtype=np.float64) - 1.0) < 1e-5
    assert np.allclose(mass, want, rtol=2e-5, atol=1e-9)
    assert (prob >= 0).all() and (prob <= 1.0).all() and (alias < n).all()
    # probability of drawing texel i = (prob[i] + sum over buckets j aliased to i of (1 - prob[j])) / n
    p = prob.astype(np.float64).ravel().copy()
    np.add.at(p, alias.ravel().astype(np.int64), 1.0 - prob.astype(np.float64).ravel())
    assert np.allclose(p / n, want.ravel(), rtol=1e-4, atol=2e-9)


def test_samples_follow_the_pdf(tpt, ctx):
    img = spot_sky()
    h, w = img.shape[:2]
    env = tpt.EnvLight(ctx, img)
    n = 1 << 21
    s = env.sample(n, seed=3)
    d, pdf = s[:, :3].astype(np.float64), s[:, 3].astype(np.float64)
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=2e-6)
    assert (pdf > 0).all()
    want = texel_weights(img)
    x, y = dir_to_texel(d, w, h)
    # the reported pdf: uniform in (u, v) inside the texel, d(omega) = 2 pi^2 sin(theta) du dv
    sin_t = np.hypot(d[:, 0], d[:, 2])
    ok = np.isclose(pdf, want[y, x] * (w * h) / (2 * np.pi ** 2 * sin_t), rtol=5e-4)
    assert ok.mean() > 0.999, ok.mean()       # the rest sit on a texel edge, where the float32 direction maps to the neighbour
    # histogram over texels against the mass: normalised chi-square ~ 1
    hist = np.bincount(y * w + x, minlength=w * h).astype(np.float64)
    exp = want.ravel() * n
    big = exp > 20
    chi2 = (((hist - exp) ** 2 / np.maximum(exp, 1e-30))[big]).sum() / big.sum()
    assert 0.8 < chi2 < 1.25, chi2
    assert hist[~big].sum() <= 1.5 * exp[~big].sum() + 50
    # Monte-Carlo integral of the map's luminance over the sphere, against the texel sum
    lum = (0.2126 * img[..., 0].astype(np.float64) + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]) / 255.0
    theta = (1.0 - (np.arange(h) + 0.5) / h) * np.pi
    exact = (lum * (2 * np.pi / w) * (2.0 * np.sin(np.pi / (2 * h)) * np.sin(theta))[:, None]).sum()   # exact band solid angles
    mc = (lum[y, x] / pdf).mean()
    assert abs(mc - exact) <= 5e-3 * exact, (mc, exact)
    # a different seed draws different samples; the same seed the same
    assert (env.sample(1024, seed=3) == s[:1024]).all()
    assert (env.sample(1024, seed=4) != s[:1024]).any()


def test_samples_follow_the_pdf_on_a_4096x2048_map(tpt, ctx):
    """bench.py's sky has 2^23 texels: bucket choice and the accept / alias coin must come from independent random bits (they
    used to share one 24-bit uniform, which left the coin one bit: texels with acceptance probability <= 0.5 were kept half the
    time, the others always).  The sampled histogram over 64 x 32 blocks of texels must follow the block masses, and the
    alias branch must be taken as often as the tables say."""
    img = synthetic_sky(4096, 2048)
    h, w = img.shape[:2]
    env = tpt.EnvLight(ctx, img)
    prob, alias, mass = env.sampler_tables()
    n = 1 << 22
    s = env.sample(n, seed=11)
    d = s[:, :3].astype(np.float64)
    x, y = dir_to_texel(d, w, h)
    bw, bh = w // 64, h // 32
    blk = (y // bh) * 64 + (x // bw)
    hist = np.bincount(blk, minlength=64 * 32).astype(np.float64)
    exp = mass.astype(np.float64).reshape(32, bh, 64, bw).sum(axis=(1, 3)).ravel() * n
    big = exp > 50
    chi2 = (((hist - exp) ** 2 / np.maximum(exp, 1e-30))[big]).sum() / big.sum()
    assert 0.8 < chi2 < 1.3, chi2
    # the sun disc (texels of value 255 in all channels) holds a known share of the mass
    sun = (img[..., :3] == 255).all(axis=-1)
    share = mass.astype(np.float64)[sun].sum()
    got = sun[y, x].mean()
    assert abs(got - share) <= 4.0 * np.sqrt(share * (1 - share) / n) + 1e-4, (got, share)
    # Monte-Carlo integral of the luminance against the texel sum (a biased texel choice shows here at once)
    lum = (0.2126 * img[..., 0].astype(np.float64) + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]) / 255.0
    theta = (1.0 - (np.arange(h) + 0.5) / h) * np.pi
    exact = (lum * (2 * np.pi / w) * (2.0 * np.sin(np.pi / (2 * h)) * np.sin(theta))[:, None]).sum()
    mc = (lum[y, x] / s[:, 3].astype(np.float64)).mean()
    assert abs(mc - exact) <= 3e-3 * exact, (mc, exact)


def test_nee_mean_matches_the_default_estimator_under_the_bench_sky(tpt, ctx, built, scenes):
    """TPT_RENDER_ENV_NEE under the 4096 x 2048 sky bench.py uses: same expectation as the default estimator."""
    ds, _ = built["ball"]
    cam = tpt.Camera.from_scene(scenes["ball"])
    W, H = 64, 36
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=synthetic_sky(4096, 2048))

    def img(spp, seed, nee):
        pt.m_radiance.clear()
        pt.render_into(ds, cam, spp=spp, max_depth=8, seed=seed, env_nee=nee)
        return pt.m_radiance.to_f32() / spp

    ref = img(1 << 14, 5, False)
    nee = img(1 << 14, 6, True)
    assert _relmse(nee, ref) <= 1e-3
    assert abs(nee.mean() - ref.mean()) <= 5e-3 * ref.mean(), (nee.mean(), ref.mean())


def test_env_sample_needs_a_map(tpt, ctx):
    env = tpt.EnvLight(ctx, None)
    with pytest.raises(RuntimeError, match="no environment map"):
        env.sample(16)


def _relmse(a, b):
    return float(((a.astype(np.float64) - b) ** 2).sum() / max((b.astype(np.float64) ** 2).sum(), 1e-30))


@pytest.mark.parametrize("name", ["ball", "box2"])
def test_nee_keeps_the_expectation_and_lowers_variance(name, tpt, ctx, built, scenes):
    ds, _ = built[name]
    cam = tpt.Camera.from_scene(scenes[name])
    W, H = 64, 36
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=spot_sky())

    def img(spp, seed, nee):
        pt.m_radiance.clear()
        pt.render_into(ds, cam, spp=spp, max_depth=8, seed=seed, env_nee=nee)
        return pt.m_radiance.to_f32() / spp

    ref = img(1 << 16, 5, False)                 # the reference's estimator, converged
    nee = img(1 << 14, 6, True)
    assert np.isfinite(nee).all()
    err = _relmse(nee, ref)
    assert err <= 1e-3, f"{name}: NEE image differs from the default estimator's: rel-MSE {err:.3e}"
    assert abs(nee.mean() - ref.mean()) <= 0.01 * ref.mean()
    # variance at equal (low) spp
    e_plain = _relmse(img(64, 7, False), ref)
    e_nee = _relmse(img(64, 7, True), ref)
    # ball is lit by the map alone; box2's noise comes from its interior bounces, which NEE does not touch
    assert e_nee < (0.5 if name == "ball" else 1.1) * e_plain, (e_nee, e_plain)
    c = ctx.counters()
    assert c["queue_overflow"] == 0


def test_nee_flag_without_a_map_changes_nothing(tpt, ctx, built, scenes):
    ds, _ = built["box"]
    cam = tpt.Camera.from_scene(scenes["box"])
    pt = tpt.PathTracer(ctx, 64, 36)
    pt.render_into(ds, cam, spp=8, seed=2)
    a = pt.m_radiance.to_i64()
    pt.m_radiance.clear()
    pt.render_into(ds, cam, spp=8, seed=2, env_nee=True)
    assert (pt.m_radiance.to_i64() == a).all()


def test_default_render_ignores_sampler_state(tpt, ctx, built, scenes):
    """Building the alias tables (or having rendered with NEE) must not change a default render's bits."""
    ds, _ = built["ball"]
    cam = tpt.Camera.from_scene(scenes["ball"])
    pt = tpt.PathTracer(ctx, 64, 36, env_rgba8=synthetic_sky(128, 64))
    pt.render_into(ds, cam, spp=8, seed=2)
    a = pt.m_radiance.to_i64()
    pt.m_radiance.clear()
    pt.render_into(ds, cam, spp=8, seed=2, env_nee=True)
    pt.m_radiance.clear()
    pt.render_into(ds, cam, spp=8, seed=2)
    assert (pt.m_radiance.to_i64() == a).all()


# ---- HDR (float) environment maps: tpt_env_set_rgb32f, SURVEY §8f-2 ---------------------------------------------------
def hdr_sky(w=128, h=64):
    """Linear float map with a sun 300x the 8-bit ceiling (what an 8-bit map cannot hold)."""
    img = np.zeros((h, w, 3), np.float32)
    img[...] = (0.05, 0.07, 0.12)
    img[h // 2:] = (0.25, 0.45, 0.9)
    img[int(h * 0.70):int(h * 0.76), int(w * 0.10):int(w * 0.13)] = (300.0, 270.0, 220.0)
    return img


def test_float_map_with_8bit_values_gives_the_8bit_maps_bits(tpt, ctx, built, scenes):
    """A float map holding exactly u8 * (1/255) (one fp32 product, what the RGBA8 lookup computes) must render the same
    accumulator bits as the RGBA8 map: the two texel formats share one lookup."""
    ds, _ = built["ball"]
    cam = tpt.Camera.from_scene(scenes["ball"])
    sky8 = synthetic_sky(128, 64)
    skyf = sky8[..., :3].astype(np.float32) * np.float32(1.0 / 255.0)
    W, H = 96, 54
    pa = tpt.PathTracer(ctx, W, H, env_rgba8=sky8)
    pa.render_into(ds, cam, spp=40, seed=4)
    a = pa.m_radiance.to_i64()
    pb = tpt.PathTracer(ctx, W, H, env_rgba8=skyf)
    pb.render_into(ds, cam, spp=40, seed=4)
    assert (pb.m_radiance.to_i64() == a).all() and a.any()
    # and switching back replaces the float map
    pc = tpt.PathTracer(ctx, W, H, env_rgba8=sky8)
    pc.render_into(ds, cam, spp=40, seed=4)
    assert (pc.m_radiance.to_i64() == a).all()


@pytest.mark.parametrize("name", ["ball", "box2"])
def test_hdr_map_render_vs_oracle(name, tpt, ctx, built, scenes, orc):
    """Same RNG on both sides (as test_render_same_rng_vs_oracle) under a float map with a 300x sun."""
    ds, _ = built[name]
    env = hdr_sky()
    W, H, spp = 96, 54, 32
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)
    pt.render_into(ds, tpt.Camera.from_scene(scenes[name]), spp=spp, max_depth=8, seed=11)
    img = pt.m_radiance.to_f32() / spp
    ref = orc.render(scenes[name], W, H, spp, max_depth=8, seed=11, env=env) / spp
    assert np.isfinite(img).all() and img.max() > 2.0          # values above the 8-bit ceiling reach the image
    err = _relmse(img, ref)
    assert err <= 2e-2, f"{name}: rel-MSE at {spp} spp with shared RNG under an HDR map = {err:.3e}"
    assert ctx.counters()["dropped_nonfinite"] == 0


def test_hdr_map_sampler_and_nee(tpt, ctx, built, scenes):
    """Alias tables are built from the float texels (mass = luminance * sin(theta)); NEE keeps the expectation under the HDR sun
    and cuts the error at equal spp."""
    env = hdr_sky()
    el = tpt.EnvLight(ctx, env)
    _, _, mass = el.sampler_tables()
    lum = 0.2126 * env[..., 0].astype(np.float64) + 0.7152 * env[..., 1] + 0.0722 * env[..., 2]
    theta = (1.0 - (np.arange(env.shape[0]) + 0.5) / env.shape[0]) * np.pi
    want = lum * np.sin(theta)[:, None]
    want /= want.sum()
    np.testing.assert_allclose(mass, want, rtol=2e-5, atol=1e-12)
    ds, _ = built["ball"]
    cam = tpt.Camera.from_scene(scenes["ball"])
    W, H = 64, 36
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)

    def img(spp, seed, nee):
        pt.m_radiance.clear()
        pt.render_into(ds, cam, spp=spp, max_depth=8, seed=seed, env_nee=nee)
        return pt.m_radiance.to_f32() / spp

    ref = img(1 << 18, 5, False)                 # the reference's estimator; the sun makes it converge slowly
    nee = img(1 << 16, 6, True)
    assert _relmse(nee, ref) <= 3e-3
    assert _relmse(img(64, 7, True), ref) < 0.5 * _relmse(img(64, 7, False), ref)


def test_float_map_rejects_bad_texels(tpt, ctx):
    bad = hdr_sky(16, 8)
    for v in (np.nan, np.inf, -1.0, 1e6):
        b = bad.copy()
        b[3, 5, 1] = v
        with pytest.raises(tpt.TptError):
            tpt.EnvLight(ctx, b)
