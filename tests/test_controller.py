"""Headless camera controller (host logic of Camera::updateCamera, reference Core/Camera.cuh:46-157) and the
progressive loop (Raytracer::run, src/Raytracer.h:52-71) driven by a script instead of SFML events."""
import importlib

import numpy as np
import pytest

SEED = 1804289383


def _cam(crt):
    cam = crt.Camera()
    cam.position[:] = (0.0, 4.0, 4.0)          # Raytracer.h:79
    cam.world_up[:] = (0.0, 1.0, 0.0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture, cam.focus_dist = -90.0, 0.0, 80.0, 16 / 9, 1e-6, 5.0
    return cam


def test_controller_motion_rules_match_reference(crt):
    ct = importlib.import_module("raytracer-cuda_b200.controller")
    c = ct.CameraController(_cam(crt), 2560, 1440)
    # still, not high quality: 1 spp
    assert c.update(1 / 60) == 1 and not c.in_motion()
    # F toggles high quality: 2000 spp while still (Camera.cuh:52-65)
    assert c.update(1 / 60, keys=("F",)) == 2000 and c.high_quality
    assert c.update(1 / 60) == 2000
    # W moves along -front (yaw -90, pitch 0: front = (0,0,1) so position.z decreases), drops to 1 spp and
    # clears high quality (Camera.cuh:57-61, :145-146)
    z0 = c.cam.position[2]
    assert c.update(0.5, keys=("W",)) == 1 and c.in_motion() and not c.high_quality
    assert abs(c.cam.position[2] - (z0 - 0.5)) < 1e-6 and abs(c.cam.position[0]) < 1e-6
    assert c.update(1 / 60) == 1 and not c.in_motion()
    # Space / LControl move along world up; A/D along right = unit(cross(front, up)) = (-1,0,0) for this pose
    y0 = c.cam.position[1]
    c.update(0.25, keys=("Space",))
    assert abs(c.cam.position[1] - (y0 + 0.25)) < 1e-6
    x0 = c.cam.position[0]
    c.update(0.25, keys=("D",))
    assert abs(c.cam.position[0] - (x0 - 0.25)) < 1e-6
    # focus: clamped at 0.1 (Camera.cuh:79-83)
    c.adjust_focus_distance(-100.0)
    assert abs(c.cam.focus_dist - 0.1) < 1e-7


def test_controller_mouse_look_smoothing_and_clamp(crt):
    ct = importlib.import_module("raytracer-cuda_b200.controller")
    c = ct.CameraController(_cam(crt), 800, 600)
    # first pressed frame only latches the position (no jump, Camera.cuh:99-108)
    c.update(1 / 60, mouse=(400, 300), mouse_pressed=True)
    assert c.cam.yaw_deg == -90.0 and not c.rotates
    # second: smoothed half-way to the new position, inverted, times sensitivity 0.2 (:112-126)
    c.update(1 / 60, mouse=(420, 300), mouse_pressed=True)
    assert c.rotates and abs(c.cam.yaw_deg - (-90.0 - 10 * 0.2)) < 1e-5 and c.samples_per_pixel == 1
    # pitch clamps to +-89 (:128)
    for _ in range(200):
        c.update(1 / 60, mouse=(420, -5000), mouse_pressed=True)
    assert abs(c.cam.pitch_deg - 89.0) < 1e-6
    c.update(1 / 60, mouse_pressed=False)
    assert not c.rotates


@pytest.mark.gpu
def test_progressive_loop_through_the_abi(crt):
    ct = importlib.import_module("raytracer-cuda_b200.controller")
    import oracle_lib as ol
    sc = crt.builtin_scene("cornell")
    W = H = 64
    r = crt.Renderer(W, H, mode=crt.MODE_AUTO)
    r.set_scene(sc)
    r.initialize(SEED, 0)
    cam = sc.camera
    c = ct.CameraController(cam, W, H)
    script = [{"dt": 1 / 60}, {"dt": 0.1, "keys": ("W",)}, {"dt": 1 / 60, "keys": ("F",)}, {"dt": 1 / 60}]
    frames = ct.run_progressive(r, c, script, max_bounces=8, high_quality_spp=6)
    assert [f[0] for f in frames] == [1, 1, 6, 6] and all(f[2] > 0 for f in frames)
    assert np.all(r.accum()[..., 3] == 6)            # every frame overwrites (CUDAKernels.h:164)
    # the frames continue the per-pixel RNG streams (CUDAKernels.h:153,165): replay on the oracle
    osc = ol.OracleScene(sc.pointer())
    cam2 = crt.builtin_scene("cornell").camera
    c2 = ct.CameraController(cam2, W, H)
    rng = ol.rng_states(SEED, 0, W, H).reshape(-1)
    last = None
    for ev, (spp, _, _) in zip(script, frames):
        c2.update(ev.get("dt"), ev.get("keys", ()))
        last = osc.render(c2.cam, W, H, spp, 8, SEED, rng_state=rng, fresh=False)
    assert (r.image_rgba8() != last[1]).any(axis=-1).mean() < 2e-3
    osc.close()
    r.close()
    sc.free()
