"""Headless camera controller + progressive frame loop.

Mirrors the HOST side of the reference's interactive loop without a window:
  * CRT::Camera::updateCamera / updateRotation / updatePosition / adjustFocusDistance
    (reference: CudaRayTracer/src/Core/Camera.cuh:46-157) — yaw/pitch FPS controller with mouse smoothing,
    W/S/A/D/Space/LControl movement, the "F" high-quality toggle, and the rule
    "camera in motion -> DEFAULT_SAMPLES_PER_PIXEL (1), still + high quality -> 2000 spp" (:57-71);
  * Raytracer::run / updateAndRender (reference: src/Raytracer.h:52-71,94-102): per frame
    updateCamera -> CUDARenderer::updateCamera -> CUDARenderer::render.
Input events come from a script (a list of dicts) instead of SFML; the render goes through the C-ABI
(Renderer.update_camera / Renderer.render). There is no window, so nothing is displayed: frames stay in the
device buffer (crt_get_image_rgba8) or are written with crt_write_ppm.

Arithmetic is done in numpy float32 in the reference's statement order so that the pose handed to
crt_set_camera matches what the reference's host code would compute.
"""
import math

import numpy as np

F32 = np.float32
DEFAULT_SAMPLES_PER_PIXEL = 1          # Camera.cuh:11
HIGH_QUALITY_SAMPLES_PER_PIXEL = 2000  # Camera.cuh:64
PI = F32(3.1415926535897932385)        # Core/Utility.cuh:9

KEYS = ("W", "S", "A", "D", "Space", "LControl", "F", "PageUp", "PageDown")


def _unit(v):
    inv = F32(1) / F32(math.sqrt(float(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])))
    return np.array([inv * v[0], inv * v[1], inv * v[2]], F32)


def _cross(u, v):
    return np.array([u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]], F32)


class CameraController:
    """State of CRT::Camera that the reference mutates on the host, frame by frame."""

    def __init__(self, camera, window_width, window_height):
        """`camera` is a crt Camera struct (pose + lens); it is updated in place by update()."""
        self.cam = camera
        self.movement_speed = F32(1.0)        # Camera.cuh:26
        self.mouse_sensitivity = F32(0.2)     # Camera.cuh:27
        self.high_quality = False
        self.moves = False
        self.rotates = False
        self.samples_per_pixel = DEFAULT_SAMPLES_PER_PIXEL
        # function-local statics of updateRotation (Camera.cuh:91-96)
        self._first_mouse = True
        self._last_x = F32(window_width / 2.0)
        self._last_y = F32(window_height / 2.0)
        self._smooth_x = self._last_x
        self._smooth_y = self._last_y

    # -- derived vectors (Camera.cuh:161-169), needed for movement ------------------------------
    def _front_right(self):
        yaw, pitch = F32(self.cam.yaw_deg), F32(self.cam.pitch_deg)
        front = np.array([-F32(math.cos(yaw * PI / F32(180))) * F32(math.cos(pitch * PI / F32(180))),
                          -F32(math.sin(pitch * PI / F32(180))),
                          -F32(math.sin(yaw * PI / F32(180))) * F32(math.cos(pitch * PI / F32(180)))], F32)
        front = _unit(front)
        right = _unit(_cross(front, np.array(list(self.cam.world_up), F32)))
        return front, right

    def _update_rotation(self, mouse_x, mouse_y, mouse_pressed):   # Camera.cuh:89-138
        smooth = F32(0.5)
        if mouse_pressed:
            if self._first_mouse:
                self._last_x = self._smooth_x = F32(mouse_x)
                self._last_y = self._smooth_y = F32(mouse_y)
                self._first_mouse = False
                return                                   # first frame skipped (no jump); m_CameraRotates unchanged
            self.rotates = True
            self._smooth_x = self._smooth_x * (F32(1) - smooth) + F32(mouse_x) * smooth
            self._smooth_y = self._smooth_y * (F32(1) - smooth) + F32(mouse_y) * smooth
            xoff = (self._smooth_x - self._last_x) * -self.mouse_sensitivity
            yoff = (self._smooth_y - self._last_y) * -self.mouse_sensitivity
            self._last_x, self._last_y = self._smooth_x, self._smooth_y
            self.cam.yaw_deg = float(F32(self.cam.yaw_deg) + xoff)
            self.cam.pitch_deg = float(max(F32(-89.0), min(F32(89.0), F32(self.cam.pitch_deg) + yoff)))
        else:
            self._first_mouse = True
            self.rotates = False

    def _update_position(self, dt, keys):                          # Camera.cuh:139-157
        velocity = self.movement_speed * F32(dt)
        # NB the reference moves along m_Front / m_Right of the PREVIOUS frame (updateCameraVectors runs after
        # updatePosition, Camera.cuh:48-50) but after this frame's rotation changed yaw/pitch; m_Front is only
        # refreshed by updateCameraVectors, so the previous frame's vectors are used. We keep them cached.
        front, right = self._cached_front_right
        pos = np.array(list(self.cam.position), F32)
        prev = pos.copy()
        up = np.array(list(self.cam.world_up), F32)
        if "W" in keys: pos = pos - front * velocity
        if "S" in keys: pos = pos + front * velocity
        if "A" in keys: pos = pos - right * velocity
        if "D" in keys: pos = pos + right * velocity
        if "Space" in keys: pos = pos + up * velocity
        if "LControl" in keys: pos = pos - up * velocity
        self.moves = bool((pos != prev).any())
        self.cam.position[:] = [float(x) for x in pos]

    def update(self, dt, keys=(), mouse=(0.0, 0.0), mouse_pressed=False):
        """Camera::updateCamera (Camera.cuh:46-72). Returns the samples per pixel of this frame."""
        if not hasattr(self, "_cached_front_right"):
            self._cached_front_right = self._front_right()
        self._update_rotation(mouse[0], mouse[1], mouse_pressed)
        self._update_position(dt, set(keys))
        self._cached_front_right = self._front_right()               # updateCameraVectors
        if "F" in keys:
            self.high_quality = not self.high_quality
        if self.rotates or self.moves:
            self.samples_per_pixel = DEFAULT_SAMPLES_PER_PIXEL
            self.high_quality = False
        elif self.high_quality:
            self.samples_per_pixel = HIGH_QUALITY_SAMPLES_PER_PIXEL
        else:
            self.samples_per_pixel = DEFAULT_SAMPLES_PER_PIXEL
        return self.samples_per_pixel

    def in_motion(self):                                              # Camera.cuh:74-77
        return self.rotates or self.moves

    def adjust_focus_distance(self, delta):                           # Camera.cuh:79-83 (PgUp/PgDn, WindowManager.h:62-71)
        self.cam.focus_dist = float(max(F32(0.1), F32(self.cam.focus_dist) + F32(delta)))


def run_progressive(renderer, controller, script, max_bounces=20, high_quality_spp=None, on_frame=None):
    """Raytracer::run without the window: for every scripted frame update the camera, upload it, render.

    script: iterable of dicts {dt, keys, mouse, mouse_pressed, focus_delta}. Like the reference every frame
    OVERWRITES the image (CUDAKernels.h:164) and continues the per-pixel RNG streams (:153,165).
    Returns a list of (spp, render_ms, rays) per frame.
    """
    out = []
    for ev in script:
        if ev.get("focus_delta"):
            controller.adjust_focus_distance(ev["focus_delta"])
        spp = controller.update(ev.get("dt", 1.0 / 60.0), ev.get("keys", ()), ev.get("mouse", (0.0, 0.0)), ev.get("mouse_pressed", False))
        if spp == HIGH_QUALITY_SAMPLES_PER_PIXEL and high_quality_spp is not None:
            spp = high_quality_spp
        renderer.update_camera(controller.cam)
        st = renderer.render(spp, max_bounces)
        out.append((spp, st.render_ms, int(st.rays)))
        if on_frame is not None:
            on_frame(len(out) - 1, renderer)
    return out
