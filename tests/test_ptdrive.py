"""ptdrive: the reference's main()/display() re-hosted as a headless C++ program over the two C ABIs,
and the PNG writer of the display stage."""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from opengl_ray_tracing_b200 import scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PTDRIVE = os.path.join(ROOT, "opengl_ray_tracing_b200", "ptdrive")


def decode_png(path):
    """Minimal decoder for what pts_write_png emits (8-bit RGB/RGBA, filter 0); checks every chunk CRC."""
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, hdr, names = 8, b"", None, []
    while pos < len(b):
        n, = struct.unpack(">I", b[pos:pos + 4])
        t, d = b[pos + 4:pos + 8], b[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(t + d) == crc, "chunk CRC"
        names.append(t)
        if t == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", d)
        if t == b"IDAT":
            idat += d
        pos += 12 + n
    assert names[0] == b"IHDR" and names[-1] == b"IEND"
    w, h, depth, ctype, comp, filt, inter = hdr
    assert (depth, comp, filt, inter) == (8, 0, 0, 0) and ctype in (2, 6)
    c = 3 if ctype == 2 else 4
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * c + 1)     # zlib verifies the Adler-32
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, c)


@pytest.mark.parametrize("shape", [(1, 1, 3), (5, 7, 3), (3, 2, 4), (300, 301, 4)])     # the last one spans several 64 KiB stored blocks
def test_png_writer_round_trip(tmp_path, shape):
    img = np.random.default_rng(shape[0]).integers(0, 256, shape, dtype=np.uint8)
    p = str(tmp_path / "a.png")
    S.writePng(p, img)
    assert (decode_png(p) == img).all()


def test_png_writer_rejects_bad_input(tmp_path):
    with pytest.raises(ValueError):
        S.writePng(str(tmp_path / "b.png"), np.zeros((4, 4, 2), np.uint8))
    with pytest.raises(Exception):
        S.writePng(str(tmp_path / "no_such_dir" / "b.png"), np.zeros((4, 4, 3), np.uint8))


def test_ptdrive_is_built_and_fails_loudly_without_a_gpu(tmp_path):
    assert os.path.exists(PTDRIVE), "python __graft_entry__.py build"
    r = subprocess.run([PTDRIVE, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--variant" in r.stdout
    r = subprocess.run([PTDRIVE, "--variant", "7"], capture_output=True, text=True)
    assert r.returncode == 1 and "variant" in r.stderr
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:                       # no CPU fallback: the scene is built on the host, then pt_create must refuse
        r = subprocess.run([PTDRIVE, "--icosphere", "4", "--size", "32x24", "--spp", "1", "--out", str(tmp_path / "x.png")],
                           capture_output=True, text=True)
        assert r.returncode == 1 and "pt_create" in r.stderr and not os.path.exists(tmp_path / "x.png")


@pytest.mark.gpu
@pytest.mark.parametrize("variant,kind", [(2, "is"), (0, "lambert"), (1, "disney")])
def test_ptdrive_image_equals_the_python_mirror(tmp_path, variant, kind):
    from opengl_ray_tracing_b200 import renderer as R, scenes
    W, H, SPP = 160, 96, 6
    png, pfm = str(tmp_path / "o.png"), str(tmp_path / "o.pfm")
    r = subprocess.run([PTDRIVE, "--variant", str(variant), "--icosphere", "22", "--size", "%dx%d" % (W, H), "--spp", str(SPP),
                        "--spp-per-call", "4", "--synth-hdr", "512x256", "--camera", "10,25,2.5", "--out", png, "--pfm", pfm],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Mrays/s" in r.stdout
    sc = scenes.build(kind, "10k", hdr_res=(512, 256), camera=(10.0, 25.0, 2.5))
    pt = R.PathTracer(W, H, sc.variant)
    pt.upload_scene(sc.tris, sc.nodes, sc.hdr, sc.hdrCache)
    pt.set_camera(sc.eye, sc.cam)
    pt.render_spp(0, SPP)
    want8, want = pt.readback_display(), pt.readback()
    pt.close()
    got8 = decode_png(png)
    assert got8.shape == (H, W, 4) and (got8 == want8).all()
    with open(pfm, "rb") as f:                                  # PF, "W H", "-1.0" (little endian), rows bottom first
        assert f.readline() == b"PF\n" and f.readline() == b"%d %d\n" % (W, H) and f.readline() == b"-1.0\n"
        rad = np.frombuffer(f.read(), "<f4").reshape(H, W, 3)
    assert rad.tobytes() == np.ascontiguousarray(want[..., :3]).tobytes()


@pytest.mark.gpu
def test_ptdrive_basic_writes_the_reference_tracers_image(tmp_path):
    from opengl_ray_tracing_b200 import basic as B
    png = str(tmp_path / "basic.png")
    r = subprocess.run([PTDRIVE, "--basic", "--size", "64x48", "--spp", "5", "--seed", "9", "--out", png], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    img, _ = B.render(B.default_scene(), 64, 48, 5, 9)
    assert (decode_png(png) == B.imshow(img)).all()
