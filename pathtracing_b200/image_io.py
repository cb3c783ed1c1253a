"""Image file output for the headless consumers: the reference's only sink is a
swapchain (RayTracing/Graphics/Renderer.cs:976-991), which cannot exist here, so frames end in a
host buffer and, when wanted, a file.  Host-side byte shuffling only; no pixel arithmetic."""
from __future__ import annotations

import numpy as np


def write_ppm(path: str, rgba8: np.ndarray) -> None:
    """Binary PPM (P6) from an (H, W, 4) RGBA8 array; alpha is dropped."""
    a = np.ascontiguousarray(rgba8, np.uint8)
    h, w = a.shape[:2]
    with open(path, "wb") as f:
        f.write(f"P6\n{w} {h}\n255\n".encode())
        f.write(a[..., :3].tobytes())


def read_ppm(path: str) -> np.ndarray:
    """(H, W, 3) uint8 from a binary PPM written by write_ppm / host/pt_render."""
    with open(path, "rb") as f:
        data = f.read()
    parts = data.split(b"\n", 3)
    assert parts[0] == b"P6" and parts[2] == b"255"
    w, h = (int(v) for v in parts[1].split())
    return np.frombuffer(parts[3], np.uint8, w * h * 3).reshape(h, w, 3)


def write_png(path: str, rgba8: np.ndarray, bgra: bool = False) -> None:
    """PNG via Pillow from RGBA8 (or the BGRA8 present target)."""
    from PIL import Image
    a = np.ascontiguousarray(rgba8, np.uint8)
    if bgra:
        a = a[..., [2, 1, 0, 3]]
    Image.fromarray(a, "RGBA").save(path)
