"""gi_debug_view: the reference's `visualize` debug views (src/visualize.cpp:268-298, 436-530, 589-700, 1366-1424) as CUDA
splat / line kernels.  The reference draws them with OpenGL, which this machine cannot run, so the checks are geometric:
a photon of known position and colour must land on the pixel the scene camera projects it to (the inverse of the
eye-ray generator, render.cpp:64-117), geometry must hide what lies behind it, the incident-direction span must point the
way visualize.cpp draws it, and the ray / path views must mark the pixels their grid rays go through."""
import os
import subprocess
import numpy as np
import pytest

import gi_b200
from gi_testlib import GOLDEN, GI_PHOTON, PKG, read_png_rgb

pytestmark = pytest.mark.gpu
W = H = 200


def _project(cam, p, w, h):
    """pixel (x, y from the bottom) of world point p; yfov follows the frame's aspect ratio (R3Viewer.cpp:76-84)"""
    eye, tow, up, right = (np.array(list(v)) for v in (cam.eye, cam.towards, cam.up, cam.right))
    v = np.asarray(p) - eye
    z = v @ tow
    tx = np.tan(cam.xfov); ty = (h / w) * tx
    sx = ((v @ right) / (z * tx) * w + 2 * (w // 2)) * 0.5
    sy = ((v @ up) / (z * ty) * h + 2 * (h // 2)) * 0.5
    return int(np.floor(sx)), int(np.floor(sy)), z


def _photon(pos, rgbe, phi, theta):
    ph = np.zeros(1, dtype=GI_PHOTON)
    ph["position"] = pos; ph["rgbe"] = rgbe; ph["direction"] = phi * 256 + theta
    return ph


@pytest.fixture(scope="module")
def ctx():
    c = gi_b200.Context(0)
    c.upload_scene(gi_b200.HostScene(os.path.join(GOLDEN, "cornell.gisc")))
    yield c
    c.close()


def test_photons_land_where_the_camera_projects_them(ctx):
    scene = gi_b200.HostScene(os.path.join(GOLDEN, "cornell.gisc"))
    cam = scene.view.camera
    eye, tow, up, right = (np.array(list(v)) for v in (cam.eye, cam.towards, cam.up, cam.right))
    # three photons floating in front of the camera (nothing of the scene before them), pure red / green / blue,
    # incident direction code (phi 128, theta 128): about +x in the direction table (photon_utils.cpp:253-272)
    pts = [eye + tow * 1.0 + right * 0.05 * k + up * 0.04 * k for k in (-1, 0, 1)]
    cols = [(255, 0, 0), (0, 255, 0), (0, 0, 255)]
    ph = np.concatenate([_photon(p, (*c, 128), 128, 128) for p, c in zip(pts, cols)])
    tp, tt = 128 * 2 * np.pi / 255.0 - np.pi, 128 * np.pi / 255.0
    incident = np.array([np.sin(tt) * np.cos(tp), np.sin(tt) * np.sin(tp), np.cos(tt)])
    ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
    ctx.build_maps()
    img = ctx.debug_view(gi_b200.GI_VIEW_CAUSTIC_PHOTONS, W, H, draw_shapes=False)
    assert img.shape == (H, W, 3)
    for p, c in zip(pts, cols):
        x, y, _ = _project(cam, p, W, H)
        assert tuple(img[y, x]) == c, (x, y, img[y, x])
        # the sphere of radius 0.01 at depth 1 covers radius / (depth tan xfov) * W / 2 pixels around it
        r = int(0.01 / (1.0 * np.tan(cam.xfov)) * W * 0.5)
        assert r >= 1 and tuple(img[y + r, x]) == c and tuple(img[y, x - r]) == c
    assert (img.reshape(-1, 3).sum(1) > 0).sum() < 0.02 * W * H          # black elsewhere (no shapes drawn)
    # the caustic view draws +10 radius along the incident direction (visualize.cpp:1406, 1421): the far end of the span
    end = pts[1] + incident * 0.1
    ex, ey, _ = _project(cam, end, W, H)
    assert tuple(img[ey, ex]) == cols[1]
    # the global view draws -2 radius (visualize.cpp:1372, 1393): same photons loaded as the global map
    ctx.load_photons(gi_b200.GI_MAP_GLOBAL, ph)
    ctx.build_maps()
    img_g = ctx.debug_view(gi_b200.GI_VIEW_GLOBAL_PHOTONS, W, H, draw_shapes=False)
    x, y, _ = _project(cam, pts[1], W, H)
    assert tuple(img_g[y, x]) == cols[1]
    assert tuple(img_g[ey, ex]) != cols[1]                               # the long caustic span is not there


def test_geometry_hides_photons_behind_it(ctx):
    scene = gi_b200.HostScene(os.path.join(GOLDEN, "cornell.gisc"))
    cam = scene.view.camera
    eye, tow = np.array(list(cam.eye)), np.array(list(cam.towards))
    # the centre eye ray's first hit; one photon just in front of it, one well behind it
    hit = ctx.intersect(np.concatenate([eye, tow])[None, :])[0]
    assert hit["hit"] == 1
    t = float(hit["t"])
    front, behind = eye + tow * (t - 0.05), eye + tow * (t + 0.5)
    for pos, seen in ((front, True), (behind, False)):
        ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, _photon(pos, (255, 0, 255, 128), 0, 128))
        ctx.build_maps()
        img = ctx.debug_view(gi_b200.GI_VIEW_CAUSTIC_PHOTONS, W, H, draw_shapes=True)
        x, y, _ = _project(cam, pos, W, H)
        assert (tuple(img[y, x]) == (255, 0, 255)) == seen
        grey = img[y, x] if not seen else img[y + 30, x]
        assert grey[0] == grey[1] == grey[2]                            # the shape pass is a grey head-light shade


def test_ray_and_path_views(ctx):
    scene = gi_b200.HostScene(os.path.join(GOLDEN, "cornell.gisc"))
    cam = scene.view.camera
    # DrawRays: 20 x 20 grid, ray (i, j) = (istep / 2 + a istep, jstep / 2 + b jstep); a white sphere at every hit
    img = ctx.debug_view(gi_b200.GI_VIEW_RAYS, W, H, draw_shapes=False)
    step = W // 20
    hits = 0
    for a in range(20):
        for b in range(20):
            i, j = step // 2 + a * step, step // 2 + b * step
            if tuple(img[j, i]) == (255, 255, 255):
                hits += 1
    assert hits > 300                                                   # the box fills the view
    # DrawPaths: only eye rays that meet a specular / transparent surface are followed (the two spheres of cornell.scn);
    # their first segment starts at the camera, so it projects to the ray's own pixel, in the pixel's colour code
    img = ctx.debug_view(gi_b200.GI_VIEW_PATHS, W, H, draw_shapes=False)
    lit = img.reshape(-1, 3).sum(1) > 0
    assert 50 < lit.sum() < 0.5 * W * H
    assert (img.reshape(-1, 3)[lit] == 255).all(1).any()                # white vertex spheres
    assert ((img.reshape(-1, 3)[lit] != 255).any(1)).any()              # coloured spans


def test_cli_writes_the_debug_views(tmp_path):
    exe = os.path.join(PKG, "bin", "photonmap")
    out = tmp_path / "o.png"
    views = {k: tmp_path / f"{k}.png" for k in ("global", "caustic", "rays", "paths")}
    flags = "-resolution 64 64 -aa 0 -global 2000 -caustic 5000 -it 2 -st 1 -tt 1 -lt 2 -ss 2 -md 2 -seed 3".split()
    for k, p in views.items():
        flags += [f"-viz_{k}", str(p)]
    r = subprocess.run([exe, os.path.join(GOLDEN, "cornell.gisc"), str(out), *flags, "-v"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert r.stdout.count("Wrote debug view to") == 4
    imgs = {k: read_png_rgb(p) for k, p in views.items()}
    for k, im in imgs.items():
        assert im.shape == (64, 64, 3) and im.std() > 0
    assert not np.array_equal(imgs["global"], imgs["caustic"])
