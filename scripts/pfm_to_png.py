#!/usr/bin/env python
"""PFM (as written by mrt_render, already tone-mapped / sRGB-encoded with --tone --srgb) -> 8-bit PNG, for looking at.
usage: pfm_to_png.py in.pfm out.png"""
import sys
import numpy as np
import cv2

def read_pfm(path):
    with open(path, "rb") as f:
        kind = f.readline().strip()
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        ch = 3 if kind == b"PF" else 1
        a = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4").reshape(h, w, ch)
    return a[::-1]                                   # PFM rows run bottom to top

img = np.nan_to_num(read_pfm(sys.argv[1]), nan=0.0, posinf=1.0, neginf=0.0)
cv2.imwrite(sys.argv[2], (np.clip(img[..., ::-1], 0, 1) * 255 + 0.5).astype(np.uint8))
