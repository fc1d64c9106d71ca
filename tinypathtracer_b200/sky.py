"""Procedural environment map used where the reference's kloppenheim_07_puresky.jpg (a missing blob) is named."""
import numpy as np


def synthetic_sky(width: int = 1024, height: int = 512) -> np.ndarray:
    """SURVEY §8(d)-2 synthetic sky (replaces the missing kloppenheim_07_puresky.jpg): RGBA8, row 0 = BOTTOM."""
    v = (np.arange(height, dtype=np.float64) + 0.5) / height          # v = 1 - theta/pi ; row 0 bottom
    u = (np.arange(width, dtype=np.float64) + 0.5) / width
    theta = (1.0 - v) * np.pi
    phi = u * 2.0 * np.pi
    dy = np.cos(theta)[:, None] * np.ones((1, width))
    st = np.sin(theta)[:, None]
    dx = st * np.cos(phi)[None, :]
    dz = st * np.sin(phi)[None, :]
    t = np.sqrt(np.maximum(dy, 0.0))
    horizon = np.array([0.9, 0.9, 0.95])
    zenith = np.array([0.25, 0.45, 0.9])
    rgb = horizon[None, None, :] * (1 - t[..., None]) + zenith[None, None, :] * t[..., None]
    rgb = np.where((dy < 0)[..., None], np.array([0.35, 0.33, 0.3])[None, None, :], rgb)
    el, az = np.radians(40.0), np.radians(30.0)
    sun = np.array([np.cos(el) * np.cos(az), np.sin(el), np.cos(el) * np.sin(az)])
    cosang = dx * sun[0] + dy * sun[1] + dz * sun[2]
    rgb = np.where((cosang > np.cos(np.radians(3.0)))[..., None], 1.0, rgb)
    img = np.zeros((height, width, 4), np.uint8)
    img[..., :3] = np.clip(np.floor(rgb * 255.0 + 0.5), 0, 255).astype(np.uint8)
    img[..., 3] = 255
    return img
