"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's JPEG defense
(`_jpeg_compress_batch`, defense_experiments.py:188-207): clamp -> ToPILImage (x*255 truncated to uint8) ->
PIL `save(format="JPEG", quality=q)` -> `Image.open(...).convert("RGB")` -> ToTensor (/255) -> clamp.

The arithmetic lives in a third-party dependency that is not vendored under /root/reference: PIL's bundled
libjpeg-turbo (here: Pillow 12.2.0, libjpeg-turbo 3.x, JPEG API 6.2; the reference pins no version).  The entropy
coding is lossless, so the round trip is the published baseline-JPEG sample pipeline of libjpeg, restated below in
numpy integer arithmetic (function names of the libjpeg sources in each docstring):

  encoder  RGB->YCbCr (jccolor.c rgb_ycc_convert) -> edge replication (jcprepct.c / jcsample.c expand_*_edge)
           -> 2x2 chroma box filter with alternating bias (jcsample.c h2v2_downsample) -> level shift
           -> 8x8 forward DCT, slow-but-accurate integer method (jfdctint.c) -> quantisation, round half away
           from zero, against the Annex-K tables scaled for the quality (jcparam.c jpeg_set_quality, force_baseline)
  decoder  dequantise -> inverse DCT (jidctint.c) + range limit -> "fancy" triangle chroma upsampling
           (jdsample.c h2v2_fancy_upsample, edge rows/columns replicated) -> YCbCr->RGB (jdcolor.c)

Pinned: tests/test_oracle_golden.py checks this restatement bit for bit against vectors produced by the reference's
own function running on PIL (oracle/gen_golden.py -> tests/golden/jpeg_q75.npz), several sizes and qualities."""
from __future__ import annotations

import numpy as np

_STD_LUMA = np.array([
    16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56,
    14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99], dtype=np.int64).reshape(8, 8)
_STD_CHROMA = np.array([
    17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99, 99,
    47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99], dtype=np.int64).reshape(8, 8)


def quant_tables(quality: int):
    """jcparam.c jpeg_quality_scaling + jpeg_add_quant_table(force_baseline=TRUE)."""
    q = min(max(int(quality), 1), 100)
    scale = 5000 // q if q < 50 else 200 - 2 * q
    out = []
    for base in (_STD_LUMA, _STD_CHROMA):
        t = (base * scale + 50) // 100
        out.append(np.clip(t, 1, 255).astype(np.int64))
    return out[0], out[1]


def _fix(x: float) -> int:
    return int(x * 65536 + 0.5)


def rgb_to_ycc(rgb: np.ndarray):
    """jccolor.c rgb_ycc_convert: 16-bit fixed point, chroma offset 128 with the ONE_HALF-1 rounding fudge."""
    r, g, b = (rgb[..., i].astype(np.int64) for i in range(3))
    half = 1 << 15
    off = 128 << 16
    y = (_fix(0.29900) * r + _fix(0.58700) * g + _fix(0.11400) * b + half) >> 16
    cb = (-_fix(0.16874) * r - _fix(0.33126) * g + _fix(0.50000) * b + off + half - 1) >> 16
    cr = (_fix(0.50000) * r - _fix(0.41869) * g - _fix(0.08131) * b + off + half - 1) >> 16
    return y, cb, cr


def ycc_to_rgb(y, cb, cr) -> np.ndarray:
    """jdcolor.c ycc_rgb_convert (table construction of build_ycc_rgb_table folded in)."""
    y, cb, cr = (a.astype(np.int64) for a in (y, cb, cr))
    half = 1 << 15
    xb, xr = cb - 128, cr - 128
    r = y + ((_fix(1.40200) * xr + half) >> 16)
    b = y + ((_fix(1.77200) * xb + half) >> 16)
    g = y + ((-_fix(0.34414) * xb + half - _fix(0.71414) * xr) >> 16)
    return np.clip(np.stack([r, g, b], axis=-1), 0, 255).astype(np.uint8)


_F = dict(c0298=2446, c0390=3196, c0541=4433, c0765=6270, c0899=7373, c1175=9633, c1501=12299, c1847=15137,
          c1961=16069, c2053=16819, c2562=20995, c3072=25172)


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def _fdct_pass(d, first: bool):
    """One 1-D pass of jfdctint.c over the LAST axis of d[..., 8] (int64)."""
    c = _F
    t0, t7 = d[..., 0] + d[..., 7], d[..., 0] - d[..., 7]
    t1, t6 = d[..., 1] + d[..., 6], d[..., 1] - d[..., 6]
    t2, t5 = d[..., 2] + d[..., 5], d[..., 2] - d[..., 5]
    t3, t4 = d[..., 3] + d[..., 4], d[..., 3] - d[..., 4]
    t10, t13, t11, t12 = t0 + t3, t0 - t3, t1 + t2, t1 - t2
    out = [None] * 8
    sh = 13 - 2 if first else 13 + 2
    if first:
        out[0] = (t10 + t11) << 2
        out[4] = (t10 - t11) << 2
    else:
        out[0] = _descale(t10 + t11, 2)
        out[4] = _descale(t10 - t11, 2)
    z1 = (t12 + t13) * c["c0541"]
    out[2] = _descale(z1 + t13 * c["c0765"], sh)
    out[6] = _descale(z1 - t12 * c["c1847"], sh)
    z1, z2, z3, z4 = t4 + t7, t5 + t6, t4 + t6, t5 + t7
    z5 = (z3 + z4) * c["c1175"]
    t4, t5, t6, t7 = t4 * c["c0298"], t5 * c["c2053"], t6 * c["c3072"], t7 * c["c1501"]
    z1, z2, z3, z4 = -z1 * c["c0899"], -z2 * c["c2562"], -z3 * c["c1961"] + z5, -z4 * c["c0390"] + z5
    out[7] = _descale(t4 + z1 + z3, sh)
    out[5] = _descale(t5 + z2 + z4, sh)
    out[3] = _descale(t6 + z2 + z3, sh)
    out[1] = _descale(t7 + z1 + z4, sh)
    return np.stack(out, axis=-1)


def fdct_islow(block: np.ndarray) -> np.ndarray:
    """jfdctint.c jpeg_fdct_islow on level-shifted samples block[..., 8(row), 8(col)]; output scaled by 8."""
    d = _fdct_pass(block.astype(np.int64), True)                       # rows
    return np.swapaxes(_fdct_pass(np.swapaxes(d, -1, -2), False), -1, -2)   # columns


def _idct_pass(d, first: bool):
    """One 1-D pass of jidctint.c over the LAST axis of d[..., 8]."""
    c = _F
    z2, z3 = d[..., 2], d[..., 6]
    z1 = (z2 + z3) * c["c0541"]
    t2 = z1 - z3 * c["c1847"]
    t3 = z1 + z2 * c["c0765"]
    t0 = (d[..., 0] + d[..., 4]) << 13
    t1 = (d[..., 0] - d[..., 4]) << 13
    t10, t13, t11, t12 = t0 + t3, t0 - t3, t1 + t2, t1 - t2
    t0, t1, t2, t3 = d[..., 7], d[..., 5], d[..., 3], d[..., 1]
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * c["c1175"]
    t0, t1, t2, t3 = t0 * c["c0298"], t1 * c["c2053"], t2 * c["c3072"], t3 * c["c1501"]
    z1, z2, z3, z4 = -z1 * c["c0899"], -z2 * c["c2562"], -z3 * c["c1961"] + z5, -z4 * c["c0390"] + z5
    t0, t1, t2, t3 = t0 + z1 + z3, t1 + z2 + z4, t2 + z2 + z3, t3 + z1 + z4
    sh = 13 - 2 if first else 13 + 2 + 3
    out = [_descale(t10 + t3, sh), _descale(t11 + t2, sh), _descale(t12 + t1, sh), _descale(t13 + t0, sh),
           _descale(t13 - t0, sh), _descale(t12 - t1, sh), _descale(t11 - t2, sh), _descale(t10 - t3, sh)]
    return np.stack(out, axis=-1)


def idct_islow(coef: np.ndarray) -> np.ndarray:
    """jidctint.c jpeg_idct_islow on dequantised coefficients coef[..., 8, 8] -> samples 0..255 (range_limit)."""
    w = np.swapaxes(_idct_pass(np.swapaxes(coef.astype(np.int64), -1, -2), True), -1, -2)   # columns
    return np.clip(_idct_pass(w, False) + 128, 0, 255)                                       # rows


def _blocks(plane: np.ndarray) -> np.ndarray:
    h, w = plane.shape
    return plane.reshape(h // 8, 8, w // 8, 8).transpose(0, 2, 1, 3)


def _unblocks(b: np.ndarray) -> np.ndarray:
    nh, nw = b.shape[:2]
    return b.transpose(0, 2, 1, 3).reshape(nh * 8, nw * 8)


def _codec_plane(plane: np.ndarray, qt: np.ndarray) -> np.ndarray:
    """level shift -> fdct -> quantise (jcdctmgr.c: divisor q<<3, round half away from zero) -> dequantise -> idct."""
    v = fdct_islow(_blocks(plane.astype(np.int64)) - 128)
    q8 = qt << 3
    mag = (np.abs(v) + (q8 >> 1)) // q8
    coef = np.where(v < 0, -mag, mag) * qt
    return _unblocks(idct_islow(coef))


def _pad_edge(plane: np.ndarray, h: int, w: int) -> np.ndarray:
    return np.pad(plane, ((0, h - plane.shape[0]), (0, w - plane.shape[1])), mode="edge")


def h2v2_downsample(full: np.ndarray, ch: int, cw: int) -> np.ndarray:
    """jcsample.c h2v2_downsample on a plane already padded to (2*ch, 2*cw): bias 1,2,1,2,... along a row."""
    f = full.astype(np.int64)
    bias = np.where(np.arange(cw) % 2 == 0, 1, 2)
    return (f[0::2, 0::2] + f[0::2, 1::2] + f[1::2, 0::2] + f[1::2, 1::2] + bias[None, :]) >> 2


def h2v2_fancy_upsample(c: np.ndarray) -> np.ndarray:
    """jdsample.c h2v2_fancy_upsample: vertical 3:1 blend with the nearer/farther row (edge rows replicated by the
    main controller's context pointers), then horizontal 3:1 with +8 / +7 rounding; first and last columns special."""
    c = c.astype(np.int64)
    ch, cw = c.shape
    up = np.pad(c, ((1, 1), (0, 0)), mode="edge")
    out = np.zeros((2 * ch, 2 * cw), dtype=np.int64)
    for v in range(2):
        near = up[1:-1]
        far = up[0:-2] if v == 0 else up[2:]
        s = 3 * near + far                               # thiscolsum per column
        sp = np.pad(s, ((0, 0), (1, 1)), mode="edge")
        last, nxt = sp[:, 0:-2], sp[:, 2:]
        even = (3 * s + last + 8) >> 4
        odd = (3 * s + nxt + 7) >> 4
        # with edge replication 3*s + s = 4*s: exactly the special first/last column formulas of the source
        out[v::2, 0::2] = even
        out[v::2, 1::2] = odd
    return out


def jpeg_roundtrip_u8(rgb: np.ndarray, quality: int = 75) -> np.ndarray:
    """uint8 [H,W,3] -> uint8 [H,W,3]: what PIL save(JPEG, quality) + open + convert('RGB') returns (4:2:0)."""
    h, w, _ = rgb.shape
    qy, qc = quant_tables(quality)
    y, cb, cr = rgb_to_ycc(rgb)
    hc, wc = (h + 1) // 2, (w + 1) // 2                   # real chroma size
    hy, wy = (h + 7) // 8 * 8, (w + 7) // 8 * 8           # luma padded to whole blocks
    hcp, wcp = (hc + 7) // 8 * 8, (wc + 7) // 8 * 8       # chroma padded to whole blocks
    y_dec = _codec_plane(_pad_edge(y, hy, wy), qy)[:h, :w]
    dec = []
    for p in (cb, cr):
        # rows: replicate to an even count, downsample, then replicate the last CHROMA row (jcprepct.c);
        # columns: replicate the last full-resolution pixel out to 2*wcp (jcsample.c expand_right_edge)
        full = _pad_edge(p, 2 * hc, 2 * wcp)
        small = h2v2_downsample(full, hc, wcp)
        small = _pad_edge(small, hcp, wcp)
        d = _codec_plane(small, qc)[:hc, :wc]
        dec.append(h2v2_fancy_upsample(d)[:h, :w])
    return ycc_to_rgb(y_dec, dec[0], dec[1])


def jpeg_compress_batch(x: np.ndarray, quality: int = 75) -> np.ndarray:
    """defense_experiments.py:188-207 on a float32 [N,3,H,W] batch."""
    x = np.clip(np.asarray(x, dtype=np.float32), 0.0, 1.0)
    u8 = (x * np.float32(255.0)).astype(np.float32).astype(np.uint8)          # ToPILImage: mul(255).byte()
    out = np.empty_like(x)
    for i in range(x.shape[0]):
        rt = jpeg_roundtrip_u8(np.ascontiguousarray(u8[i].transpose(1, 2, 0)), quality)
        out[i] = (rt.astype(np.float32) / np.float32(255.0)).transpose(2, 0, 1)  # ToTensor: div(255)
    return np.clip(out, 0.0, 1.0)
