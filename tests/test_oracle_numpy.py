"""An independent statement of the reference kernel in vectorised numpy float32 (every numpy float32 operation is one
IEEE round-to-nearest operation, like the frozen spec), compared bit for bit with the C oracle on whole frames.
Follows RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:6-39 and the rounding spec of SURVEY.md App. A; shares no code
with oracle/pt_oracle.c, so a transcription error in either shows up here.  Also restates the jitter RNG (pcg3d) and
the multi-sample accumulation in numpy integers / float32."""
import numpy as np
import pytest

F = np.float32


def dot3(ax, ay, az, bx, by, bz):
    s = ax * bx                     # ((x0*y0 + x1*y1) + x2*y2), one rounding per operation
    s = s + ay * by
    s = s + az * bz
    return s


def csmain_numpy(px, py, div, cam, centre, radius):
    """px, py: float32 arrays of (float)id.xy (+ jitter).  -> colour (..., 4) float32, hit mask."""
    cam = [F(v) for v in cam]
    centre = [F(v) for v in centre]
    uvx = (px / F(div[0])) * F(2) - F(1)                                   # Test.hlsl:7
    uvy = (py / F(div[1])) * F(2) - F(1)
    rz = np.full_like(uvx, F(-1))
    ln = np.sqrt(dot3(uvx, uvy, rz, uvx, uvy, rz))                         # :10 normalize = v / sqrt(dot(v,v))
    dx, dy, dz = uvx / ln, uvy / ln, rz / ln
    oc = [cam[0] - centre[0], cam[1] - centre[1], cam[2] - centre[2]]      # :16
    a = dot3(dx, dy, dz, dx, dy, dz)                                       # :17
    b = F(2) * dot3(oc[0], oc[1], oc[2], dx, dy, dz)                       # :18
    c = dot3(oc[0], oc[1], oc[2], oc[0], oc[1], oc[2]) - F(radius) * F(radius)   # :19
    disc = b * b - (F(4) * a) * c                                          # :21
    hit = disc > F(0)                                                      # :24 (strict)
    with np.errstate(invalid="ignore", divide="ignore"):
        t = ((-b) - np.sqrt(np.where(hit, disc, F(1)))) / (F(2) * a)       # :27
        pcx = (cam[0] + t * dx) - centre[0]                                # :28-29
        pcy = (cam[1] + t * dy) - centre[1]
        pcz = (cam[2] + t * dz) - centre[2]
        nl = np.sqrt(dot3(pcx, pcy, pcz, pcx, pcy, pcz))
        nx, ny, nz = pcx / nl, pcy / nl, pcz / nl
    col = np.empty(uvx.shape + (4,), np.float32)
    col[..., 0] = np.where(hit, nx * F(0.5) + F(0.5), uvx)                 # :31 / :36
    col[..., 1] = np.where(hit, ny * F(0.5) + F(0.5), uvy)
    col[..., 2] = np.where(hit, nz * F(0.5) + F(0.5), F(0))
    col[..., 3] = F(1)
    return col, hit


def unorm8_numpy(c):
    """R8G8B8A8Unorm store (Renderer.cs:124): NaN -> 0, clamp, x255, round half to even."""
    c = np.where(np.isnan(c), F(0), c)
    return np.rint(np.clip(c, F(0), F(1)) * F(255)).astype(np.uint8)


def pixel_grid(w, h):
    x = np.broadcast_to(np.arange(w, dtype=np.float32)[None, :], (h, w))
    y = np.broadcast_to(np.arange(h, dtype=np.float32)[:, None], (h, w))
    return np.ascontiguousarray(x), np.ascontiguousarray(y)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


CASES = [
    dict(w=320, h=180),
    dict(w=1920, h=1080),                                   # Test.hlsl:6 literal: divisor (1080, 1080)
    dict(w=320, h=180, div=(1080, 1080)),                   # the literal divisor on a small image: all miss
    dict(w=333, h=77),
    dict(w=256, h=144, cam=(0.3, -0.2, 1.5), centre=(0.1, 0.05, -0.25), radius=0.7),
    dict(w=200, h=120, cam=(0, 0, 0.2), radius=0.5),        # camera inside the sphere (Test.hlsl has no t > 0 test)
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"{c['w']}x{c['h']}" + ("_moved" if "cam" in c else "") + ("_div" if "div" in c else ""))
def test_c_oracle_equals_numpy_restatement(oracle, case):
    w, h = case["w"], case["h"]
    div = case.get("div", (h, h))
    cam, centre, radius = case.get("cam", (0, 0, 1)), case.get("centre", (0, 0, 0)), case.get("radius", 0.5)
    p = oracle.frame_params(w, h, uv_div=div, cam=cam, centre=centre, radius=radius)
    ref = oracle.csmain_frame(p)
    x, y = pixel_grid(w, h)
    col, hit = csmain_numpy(x, y, div, cam, centre, radius)
    assert np.array_equal(hit, ref["hit"] >= 0) and int(hit.sum()) == ref["hits"]
    assert np.array_equal(bits(col), bits(ref["f32"]))
    assert np.array_equal(unorm8_numpy(col), ref["rgba8"])


def pcg3d_numpy(x, y, z):
    """Jarzynski & Olano's pcg3d on uint32 arrays (wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        x = x * np.uint32(1664525) + np.uint32(1013904223)
        y = y * np.uint32(1664525) + np.uint32(1013904223)
        z = z * np.uint32(1664525) + np.uint32(1013904223)
        x = x + y * z
        y = y + z * x
        z = z + x * y
        x = x ^ (x >> np.uint32(16))
        y = y ^ (y >> np.uint32(16))
        z = z ^ (z >> np.uint32(16))
        x = x + y * z
        y = y + z * x
        z = z + x * y
    return x, y, z


@pytest.mark.parametrize("size,first,n,seed", [((160, 90), 0, 5, 0), ((97, 33), 7, 3, 12345)])
def test_c_oracle_accumulation_equals_numpy_restatement(oracle, size, first, n, seed):
    """n1/n2: jitter = pcg3d(pixel, sample, seed) -> (word >> 9) * 2^-23 added to (float)id; samples summed in order."""
    w, h = size
    p = oracle.frame_params(w, h)
    ref = oracle.accumulate(p, first, n, seed, True)
    x, y = pixel_grid(w, h)
    pix = (np.arange(h, dtype=np.uint32)[:, None] * np.uint32(w) + np.arange(w, dtype=np.uint32)[None, :]).astype(np.uint32)
    sr = np.zeros((h, w, 4), np.float32)
    for s in range(first, first + n):
        vx, vy, _ = pcg3d_numpy(pix, np.full_like(pix, s), np.full_like(pix, seed))
        jx = (vx >> np.uint32(9)).astype(np.float32) * F(2.0 ** -23)
        jy = (vy >> np.uint32(9)).astype(np.float32) * F(2.0 ** -23)
        col, _ = csmain_numpy(x + jx, y + jy, (h, h), (0, 0, 1), (0, 0, 0), 0.5)
        sr = sr + col
    acc = np.zeros((h, w, 4), np.float32) + sr          # one add of the launch's partial sum into the accumulator
    assert np.array_equal(bits(acc), bits(ref))
