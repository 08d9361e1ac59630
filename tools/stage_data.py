#!/usr/bin/env python
"""Stages the reference's example INPUT DATA (not code) from the read-only mount into baseline/_data/ — git-ignored, but not
gpurun-ignored, so it travels to the GPU box where /root/reference does not exist:
    data/nerf/fox      -> baseline/_data/fox            (transforms.json + the JPEGs it names; BASELINE config 3)
    data/image/albert.exr -> baseline/_data/albert/albert_rgba_f32.npy   (1024x1024 RGBA float32 as the reference's tinyexr loader
                          delivers it; decoded here with OpenCV so that the GPU box needs no EXR codec; BASELINE config 2)
Run in the build container:  OPENCV_IO_ENABLE_OPENEXR=1 python tools/stage_data.py"""
import os
import shutil
import sys

os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
dst = os.path.join(ROOT, "baseline", "_data")

fox_src, fox_dst = os.path.join(REF, "data", "nerf", "fox"), os.path.join(dst, "fox")
if os.path.isdir(fox_src) and not os.path.isdir(fox_dst):
    shutil.copytree(fox_src, fox_dst)
    print("staged", fox_dst)

exr = os.path.join(REF, "data", "image", "albert.exr")
out = os.path.join(dst, "albert", "albert_rgba_f32.npy")
if os.path.exists(exr):
    img = cv2.imread(exr, cv2.IMREAD_UNCHANGED)          # BGR(A) float32
    assert img is not None and img.dtype == np.float32, "OpenCV could not decode the EXR (OPENCV_IO_ENABLE_OPENEXR=1?)"
    if img.ndim == 2:
        img = np.repeat(img[..., None], 3, axis=2)
    if img.shape[2] == 3:
        img = np.concatenate([img, np.ones_like(img[..., :1])], axis=2)
    img = np.ascontiguousarray(img[..., [2, 1, 0, 3]])   # -> RGBA
    os.makedirs(os.path.dirname(out), exist_ok=True)
    np.save(out, img)
    print("staged", out, img.shape, img.dtype, float(img[..., :3].min()), float(img[..., :3].max()))
