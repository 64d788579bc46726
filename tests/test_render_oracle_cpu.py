"""CPU: sanity properties of the shader restatement (oracle/render_oracle.cpp) and of the product's
host-side camera set-up.  PARITY UNPINNED against the reference's GL output (no graphics stack)."""
import numpy as np

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import render, synth


def test_camera_basis_matches_glm_lookat_formulas():
    for p in (render.default_params(640, 480), render.oblique_params(512, 512, render.ISOSURFACE)):
        b = render.camera_basis(p)
        blob = oa.camera_params(tuple(p.eye), tuple(p.front), tuple(p.up), p.vfov_deg, p.width, p.height,
                                tuple(p.step_size), p.isoValue, 0)
        want = np.frombuffer(blob[:80], np.float32)
        assert np.allclose(b, want, rtol=0, atol=1e-6)
        s, u, f = b[3:6], b[6:9], b[9:12]
        assert abs(np.dot(s, u)) < 1e-6 and abs(np.dot(s, f)) < 1e-6 and abs(np.linalg.norm(u) - 1) < 1e-6


def test_empty_volume_composites_to_white_and_iso_misses():
    vol = np.zeros((8, 16, 16), np.uint8)
    p = render.default_params(48, 36, render.COMPOSITE, eye=(0, 0, -3))  # far enough to see background
    img, S, _ = oa.oracle_render(vol, render.oracle_param_blob(p), 2)
    assert S > 0
    assert np.all(img[..., :3] == 1.0)          # 1 - 0, 1 - 0, b = 255 clamped
    cube = img[..., 3] == 0.0                    # pixels on the cube: alpha 0; background: clear colour alpha 1
    assert cube.any() and (~cube).any()
    p.mode = render.ISOSURFACE
    img, _, hits = oa.oracle_render(vol, render.oracle_param_blob(p), 2, want_hits=True)
    assert np.all(img == 1.0) and np.all(hits == -1)


def test_opaque_volume_terminates_early():
    vol = np.full((8, 16, 16), 255, np.uint8)
    p = render.default_params(32, 24, render.COMPOSITE)
    img, S, _ = oa.oracle_render(vol, render.oracle_param_blob(p), 2)
    on_cube = img[..., 3] > 0.99
    assert on_cube.any()
    # s = 1: A goes 0.6, 0.84, 0.936, 0.9744, 0.98976, 0.995904 -> 6 samples then break (raycaster.frag:77)
    assert S == 6 * int(on_cube.sum())


def test_iso_hit_on_a_ramp_is_an_upward_crossing_only():
    z = np.arange(32, dtype=np.float32)[:, None, None]
    vol = np.broadcast_to(np.clip(z * 8, 0, 255), (32, 16, 16)).astype(np.uint8)   # rising along +z = along the default view
    p = render.default_params(40, 30, render.ISOSURFACE, step=(1 / 16, 1 / 16, 1 / 32))
    img, _, hits = oa.oracle_render(vol, render.oracle_param_blob(p), 2, want_hits=True)
    assert (hits >= 0).any()
    # looking from the other side the field falls along the ray: no upward crossing -> no hit
    p2 = render.default_params(40, 30, render.ISOSURFACE, eye=(0, 0, 0.75), front=(0, 0, -1), step=(1 / 16, 1 / 16, 1 / 32))
    _, _, hits2 = oa.oracle_render(vol, render.oracle_param_blob(p2), 2, want_hits=True)
    centre = hits2[10:20, 15:25]
    assert np.all(centre == -1)
