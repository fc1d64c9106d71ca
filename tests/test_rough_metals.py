"""TPT_RENDER_ROUGH_METALS: GGX reflection for metals, from the `roughness` field the reference loads
(src/mesh.cu:231) but never reads in a kernel.  The reference has nothing to compare with, so the lobe is pinned to the
analytic microfacet BRDF: under a uniform white environment a white rough-metal plane must show the BRDF's directional
albedo, computed here by plain quadrature of  D(h) G1(wo) G1(wi) / (4 cos_o)  over the hemisphere.  The CPU oracle and the
CUDA path are both held to it, and to each other with the shared RNG."""
import numpy as np
import pytest

from conftest import load_fixture_scene


def plane_scene(roughness: float, base=(1.0, 1.0, 1.0)):
    """One 100 x 100 quad in the plane y = 0 seen from the `square` fixture's camera (centre ray: cos_o = 0.301)."""
    s = dict(load_fixture_scene("square"))
    v = np.array([[-50, 0, 50], [50, 0, 50], [-50, 0, -50], [50, 0, -50]], np.float32)
    s["vertices"] = v
    s["normals"] = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    s["indices"] = np.array([0, 1, 3, 0, 3, 2], np.uint32)
    s["texcoords"] = np.zeros((4, 2), np.float32)
    s["objects"] = np.array([(0, 0)], s["objects"].dtype)
    s["vert_trans"] = np.eye(4, dtype=np.float32).reshape(1, 16)
    s["normal_trans"] = np.eye(4, dtype=np.float32).reshape(1, 16)
    m = np.zeros(1, s["materials"].dtype)
    m["base"] = base
    m["metallic"] = 1.0
    m["unused"][0, 2] = roughness          # Material::roughness
    s["materials"] = m
    return s


def ggx_albedo(cos_o: float, alpha: float, n_theta=1500, n_phi=3000) -> float:
    th = (np.arange(n_theta) + 0.5) * (0.5 * np.pi / n_theta)
    ph = (np.arange(n_phi) + 0.5) * (2 * np.pi / n_phi)
    T, P = np.meshgrid(th, ph, indexing="ij")
    wi = np.stack([np.sin(T) * np.cos(P), np.sin(T) * np.sin(P), np.cos(T)], -1)
    wo = np.array([np.sqrt(1 - cos_o ** 2), 0.0, cos_o])
    h = wi + wo
    h /= np.linalg.norm(h, axis=-1, keepdims=True)
    a2 = alpha * alpha
    D = a2 / (np.pi * (h[..., 2] ** 2 * (a2 - 1) + 1) ** 2)
    g1 = lambda c: 2 * c / (c + np.sqrt(a2 + (1 - a2) * c * c))
    f = D * g1(cos_o) * g1(wi[..., 2]) / (4 * cos_o)
    return float((f * np.sin(T)).sum() * (0.5 * np.pi / n_theta) * (2 * np.pi / n_phi))


WHITE = np.full((8, 16, 4), 255, np.uint8)
CENTRE = (16, 9, 17, 10)   # the centre pixel of a 33 x 19 frame


def centre_cos(orc, scene):
    r = orc.primary_rays(scene, 33, 19)[9 * 33 + 16]
    return float(-r[4])          # normal is +y


@pytest.mark.parametrize("roughness", [0.3, 0.6, 1.0])
def test_oracle_rough_metal_shows_the_ggx_albedo(roughness, orc):
    s = plane_scene(roughness)
    spp = 200_000
    img = orc.render(s, 33, 19, spp, max_depth=2, seed=5, env=WHITE, region=CENTRE, rough_metals=True) / spp
    got = float(img[9, 16].mean())
    want = ggx_albedo(centre_cos(orc, s), roughness ** 2)
    assert 0.3 < want < 1.0
    assert abs(got - want) <= 0.01 * want, (got, want)
    # without the flag the metal is the reference's mirror: everything is reflected into the white sky
    img = orc.render(s, 33, 19, 64, max_depth=2, seed=5, env=WHITE, region=CENTRE) / 64
    assert abs(float(img[9, 16].mean()) - 1.0) < 1e-6


def test_oracle_smooth_metal_ignores_the_flag(orc):
    s = load_fixture_scene("box")                # its metal has roughness 0
    a = orc.render(s, 48, 27, 4, seed=3)
    b = orc.render(s, 48, 27, 4, seed=3, rough_metals=True)
    assert np.array_equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("roughness", [0.3, 1.0])
def test_gpu_rough_metal_shows_the_ggx_albedo(roughness, tpt, ctx, orc):
    s = plane_scene(roughness, base=(1.0, 0.5, 0.25))
    ds = tpt.DeviceScene(ctx, s)
    tpt.BVH(ds).construct()
    W, H, spp = 33, 19, 1 << 18
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=WHITE)
    pt.render_into(ds, tpt.Camera.from_scene(s), spp=spp, max_depth=2, seed=6, rough_metals=True)
    img = pt.m_radiance.to_f32() / spp
    want = ggx_albedo(centre_cos(orc, s), roughness ** 2)
    np.testing.assert_allclose(img[9, 16], want * np.array([1.0, 0.5, 0.25]), rtol=0.01)     # Fresnel = baseColor
    assert ctx.counters()["dropped_nonfinite"] == 0
    ds.close()


@pytest.mark.gpu
def test_gpu_rough_metals_vs_oracle_and_flag_off_is_the_mirror(tpt, ctx, scenes, orc):
    """box1.gltf has a metal with roughness 1; box.gltf's metal has roughness 0 and must not change by a bit."""
    from conftest import synthetic_sky
    env = synthetic_sky(128, 64)
    W, H, spp = 96, 54, 32
    s = scenes["box1"]
    ds = tpt.DeviceScene(ctx, s)
    tpt.BVH(ds).construct()
    cam = tpt.Camera.from_scene(s)
    pt = tpt.PathTracer(ctx, W, H, env_rgba8=env)
    pt.render_into(ds, cam, spp=spp, seed=11, rough_metals=True)
    rough = pt.m_radiance.to_f32() / spp
    ref = orc.render(s, W, H, spp, seed=11, env=env, rough_metals=True) / spp
    err = float(((rough - ref) ** 2).sum() / (ref ** 2).sum())
    assert err <= 2e-2, f"rel-MSE with shared RNG, rough metals: {err:.3e}"
    pt.m_radiance.clear()
    pt.render_into(ds, cam, spp=spp, seed=11)
    mirror = pt.m_radiance.to_f32() / spp
    assert float(((rough - mirror) ** 2).sum() / (mirror ** 2).sum()) > 1e-3          # the flag does something here
    ds.close()
    s = scenes["box"]
    ds = tpt.DeviceScene(ctx, s)
    tpt.BVH(ds).construct()
    pt = tpt.PathTracer(ctx, W, H)
    pt.render_into(ds, tpt.Camera.from_scene(s), spp=spp, seed=11)
    a = pt.m_radiance.to_i64()
    pt.m_radiance.clear()
    pt.render_into(ds, tpt.Camera.from_scene(s), spp=spp, seed=11, rough_metals=True)
    assert (pt.m_radiance.to_i64() == a).all() and a.any()
    ds.close()
