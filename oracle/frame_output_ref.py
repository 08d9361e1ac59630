"""ORACLE — TEST INFRASTRUCTURE ONLY.  numpy (float32) restatement of the output stage of a rendered frame:
  accumulate_kernel           /root/reference/src/render_buffer.cu:232-260
  tonemap(x, curve)           /root/reference/src/render_buffer.cu:262-316
  tonemap(col, exposure, ...) /root/reference/src/render_buffer.cu:318-339
  overlay_image_kernel        /root/reference/src/render_buffer.cu:348-418
  tonemap_kernel              /root/reference/src/render_buffer.cu:533-565
  read_rgba (Byte / Float)    /root/reference/include/neural-graphics-primitives/common_device.cuh:885-913
  Testbed::calc_focal_length / render_screen_center / set_camera_to_training_view   /root/reference/src/testbed.cu:841-856,4372-4379
Parity unpinned by reference outputs: the reference's render buffer needs its CUDA build (unavailable here — empty tiny-cuda-nn submodule);
the curves are pinned against their published closed forms instead (tests/test_frame_output_cpu.py)."""
import numpy as np

F = np.float32
LINEAR, SRGB, VISPOSNEG = 0, 1, 2
IDENTITY, ACES, HABLE, REINHARD = 0, 1, 2, 3


def srgb_to_linear(s):
    s = np.asarray(s, F)
    return np.where(s <= F(0.04045), s / F(12.92), np.power(np.maximum((s + F(0.055)) / F(1.055), F(0)), F(2.4))).astype(F)


def linear_to_srgb(l):
    l = np.asarray(l, F)
    return np.where(l < F(0.0031308), F(12.92) * l, F(1.055) * np.power(np.maximum(l, F(0)), F(0.41666)) - F(0.055)).astype(F)


def accumulate(frame, accum, sample_count, color_space):
    """one accumulate_kernel pass: returns the new accumulate buffer ([..., 4] float32)"""
    color, tmp = np.array(frame, F), np.array(accum, F)
    n = F(sample_count)
    if color_space == VISPOSNEG:
        val = color[..., 0] - color[..., 1]
        tmp_val = ((tmp[..., 0] - tmp[..., 1]) * n + val) / (n + F(1))
        tmp[..., 0] = np.maximum(tmp_val, F(0)); tmp[..., 1] = np.maximum(-tmp_val, F(0))
    else:
        if color_space == SRGB:
            color[..., :3] = linear_to_srgb(color[..., :3])
        tmp[..., :3] = (tmp[..., :3] * n + color[..., :3]) / (n + F(1))
    tmp[..., 3] = (tmp[..., 3] * n + color[..., 3]) / (n + F(1))
    return tmp


def tonemap_curve(x, curve):
    x = np.asarray(x, F)
    if curve == IDENTITY:
        return x
    x = np.maximum(x, F(0))
    if curve == ACES:
        k0, k1, k2, k3, k4, k5 = F(0.6) * F(0.6) * F(2.51), F(0.6) * F(0.03), F(0), F(0.6) * F(0.6) * F(2.43), F(0.6) * F(0.59), F(0.14)
    elif curve == HABLE:
        A, B, C, D, E, Fc, W = F(0.15), F(0.50), F(0.10), F(0.20), F(0.02), F(0.30), F(11.2)
        k0, k1, k2, k3, k4, k5 = A * Fc - A * E, C * B * Fc - B * E, F(0), A * Fc, B * Fc, D * Fc * Fc
        nom, denom = k0 * (W * W) + k1 * W + k2, k3 * (W * W) + k4 * W + k5
        ws = denom / nom
        k0, k1, k2, k3, k4 = F(4) * k0 * ws, F(2) * k1 * ws, k2 * ws, F(4) * k3, F(2) * k4
    else:
        Y = F(0.2126) * x[..., 0] + F(0.7152) * x[..., 1] + F(0.0722) * x[..., 2]
        return (x * (F(1) / (Y + F(1)))[..., None]).astype(F)
    sq = x * x
    return ((sq * k0 + k1 * x + k2) / (k3 * sq + k4 * x + k5)).astype(F)


def tonemap(col, exposure3, curve, color_space, output_color_space):
    col = np.asarray(col, F)
    if color_space == SRGB:
        col = srgb_to_linear(col)
    col = col * np.power(F(2), np.asarray(exposure3, F))
    col = tonemap_curve(col, curve)
    if output_color_space == SRGB:
        col = linear_to_srgb(col)
    return col.astype(F)


def tonemap_pass(accum, exposure, background, color_space, output_color_space, curve, clamp_output_color=False, unmultiply_alpha=False):
    """tonemap_kernel over a whole accumulate buffer [..., 4]"""
    bg = np.array(background, F)
    if color_space != SRGB:
        bg[:3] = srgb_to_linear(bg[:3])
    color = np.array(accum, F)
    weight = (F(1) - color[..., 3]) * bg[3]
    color[..., :3] += bg[:3] * weight[..., None]
    color[..., 3] += weight
    color[..., :3] = tonemap(color[..., :3], [exposure] * 3, curve, color_space, output_color_space)
    if unmultiply_alpha:
        a = color[..., 3:4]
        color[..., :3] = np.where(a > 0, color[..., :3] / np.where(a > 0, a, F(1)), color[..., :3])
    if clamp_output_color:
        color = np.clip(color, F(0), F(1))
    return color.astype(F)


def read_rgba(image, data_type, x, y):
    """image: [h, w, 4] uint8 (data_type 0) or float32 (1); x, y integer arrays of valid pixels"""
    if data_type == 0:
        px = image[y, x].astype(F) * (F(1) / F(255))
        masked = (image[y, x, 0] == 0xFF) & (image[y, x, 1] == 0x00) & (image[y, x, 2] == 0xFF) & (image[y, x, 3] == 0x00)
        out = np.concatenate([srgb_to_linear(px[..., :3]) * px[..., 3:4], px[..., 3:4]], axis=-1).astype(F)
        out[masked] = F(-1)
        return out
    return image[y, x].astype(F)


def overlay_image(surface, alpha, exposure3, background, output_color_space, image, data_type, curve, color_space, fov_axis, zoom, screen_center=(0.5, 0.5)):
    """overlay_image_kernel over a whole surface [H, W, 4]; returns the new surface"""
    H, W = surface.shape[:2]
    ih, iw = image.shape[:2]
    res, ires = (W, H), (iw, ih)
    scale = F(ires[fov_axis]) / F(res[fov_axis])
    x, y = np.meshgrid(np.arange(W), np.arange(H))
    fx, fy = x.astype(F) + F(0.5), y.astype(F) + F(0.5)
    fx = (fx - F(W) * F(0.5)) / F(zoom) + F(screen_center[0]) * F(W)
    fy = (fy - F(H) * F(0.5)) / F(zoom) + F(screen_center[1]) * F(H)
    u = (fx - F(W) * F(0.5)) * scale + F(iw) * F(0.5)
    v = (fy - F(H) * F(0.5)) * scale + F(ih) * F(0.5)
    sx, sy = np.floor(u).astype(np.int64), np.floor(v).astype(np.int64)
    ok = (sx >= 0) & (sy >= 0) & (sx < iw) & (sy < ih)
    color = np.zeros((H, W, 4), F)
    color[ok] = read_rgba(image, data_type, sx[ok], sy[ok])
    bg = np.array(background, F)
    if color_space != SRGB:
        bg[:3] = srgb_to_linear(bg[:3])
    else:
        a = color[..., 3:4]
        color[..., :3] = np.where(a > 0, linear_to_srgb(color[..., :3] / np.where(a > 0, a, F(1))) * a, F(0))
    weight = (F(1) - color[..., 3]) * bg[3]
    color[..., :3] += bg[:3] * weight[..., None]
    color[..., 3] += weight
    color[..., :3] = tonemap(color[..., :3], exposure3, curve, color_space, output_color_space)
    return (color * F(alpha) + np.asarray(surface, F) * (F(1) - F(alpha))).astype(F)


def view_of_training_image(meta, width, height, fov_axis=1, zoom=1.0):
    """set_camera_to_training_view + calc_focal_length + render_screen_center: (camera, focal_length px, screen_center) the tracer gets"""
    res = (width, height)
    rel = [F(meta["focal_length"][k]) / F(meta["res"][fov_axis]) for k in range(2)]
    focal = [rel[k] * F(res[fov_axis]) * F(zoom) for k in range(2)]
    sc = [F(1) - F(meta["principal_point"][k]) for k in range(2)]
    centre = [(F(0.5) - sc[k]) * F(zoom) + F(0.5) for k in range(2)]
    return np.asarray(meta["xform"], F), focal, centre
