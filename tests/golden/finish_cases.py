"""Inputs of the finishing-chain parity tests: one seeded synthetic HDR frame and the
post_processor parameter sets run through the reference (tests/golden/make_golden.py finish)."""
import numpy as np

W, H = 96, 60


def hdr_image(seed=11):
    """Linear HDR frame with what the chain reacts to: smooth gradients, saturated colours, lognormal highlights far
    above the bloom threshold, a black region (luminance floor 1e-4) and hard edges (sharpening)."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:H, 0:W]
    img = np.stack([0.05 + 0.9 * x / (W - 1), 0.1 + 0.6 * y / (H - 1), 0.4 + 0.3 * np.sin(x / 7.0) * np.cos(y / 5.0)], -1)
    img *= rng.lognormal(mean=-0.5, sigma=0.6, size=(H, W, 1))
    hot = rng.random((H, W)) < 0.02
    img[hot] *= rng.lognormal(mean=3.0, sigma=0.7, size=(int(hot.sum()), 1))
    img[10:20, 30:50] = 0.0
    img[40:50, 60:80] = np.array([2.5, 0.2, 0.1])
    img[25:30, 5:25] *= np.array([0.1, 1.5, 0.2])
    return np.ascontiguousarray(img, np.float64)


CASES = {
    "default": (0, {}),
    "bloom_sharpen_aces": (0, dict(use_bloom=1, bloom_threshold=0.8, bloom_intensity=0.4, bloom_radius=5, use_sharpening=1,
                                   sharpen_amount=0.25, use_aces=1, contrast=1.2, color_balance=(1.1, 1.0, 0.9), exposure=0.8)),
    "hsv_auto_exposure": (0, dict(saturation=1.4, hue_shift=35.0, vignette_intensity=0.0, use_auto_exposure=1,
                                  exposure_compensation_stops=0.5)),
    "desaturate_red_debug": (0, dict(saturation=0.5, hue_shift=-120.0, debug_red=1, exposure=1.0)),
    "luminance_debug": (0, dict(debug_luminance=1, exposure=0.7, use_aces=1)),
    "light_pass": (1, dict(exposure=1.0, contrast=0.9)),
    "aux_pass": (2, {}),
}


# Image export: (reference render_pass enum value, post parameters) run through
# camera::save_render_pass -> PNG -> 8-bit RGB (tests/golden/make_golden.py export).
RENDER_PASS = {"RGB": 0, "Denoise": 1, "Albedo": 2, "Normals": 3, "Reflections": 4, "Refractions": 5, "Z-Depth": 6}
EXPORT_CASES = {
    "rgb_default": (0, {}),
    "rgb_bloom_sharpen_aces": (0, dict(use_bloom=1, bloom_threshold=0.8, bloom_intensity=0.4, bloom_radius=5, use_sharpening=1,
                                       sharpen_amount=0.25, use_aces=1, contrast=1.2, color_balance=(1.1, 1.0, 0.9), exposure=0.8)),
    "rgb_hsv": (0, dict(saturation=1.4, hue_shift=35.0, vignette_intensity=0.0, exposure=0.3)),
    "albedo": (2, dict(exposure=1.0, use_bloom=1)),
    "reflections": (4, dict(exposure=1.0, contrast=0.9)),
    "z_depth": (6, {}),
}
