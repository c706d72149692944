"""The C++ host side (include/crt_host.hpp: CRT::Camera / SceneManager / CUDARenderer / WindowManager / Raytracer,
mirroring the reference's host classes, src/Raytracer.h, src/CUDARenderer.cuh, src/Core/Camera.cuh) and its
application raytracer-cuda_b200/crt_cli (the reference's EntryPoint.cu without the window)."""
import importlib
import json
import os
import random
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "raytracer-cuda_b200")
CLI = os.path.join(PKG, "crt_cli")
SEED = 1804289383


@pytest.fixture(scope="module")
def cli():
    subprocess.check_call(["make", "-C", PKG, "-j4"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    assert os.path.exists(CLI), "crt_cli was not built"
    return CLI


def test_cli_help(cli):
    out = subprocess.run([cli, "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "usage: crt_cli" in out.stdout


def test_cli_has_no_cpu_fallback(cli, gpu_available):
    """Without a CUDA device the C++ host fails like the reference's CUDA_CHECK path (exception caught in main,
    EntryPoint.cu:38-41): message on stderr, non-zero exit, no image."""
    if gpu_available:
        pytest.skip("a CUDA device is present")
    out = subprocess.run([cli, "--scene", "cornell", "--spp", "1", "--width", "16", "--height", "16"], capture_output=True, text=True)
    assert out.returncode == 1
    assert "Error: CUDA error" in out.stderr and "no CPU fallback" in out.stderr
    assert out.stdout.strip() == ""


def _script(n, seed):
    rnd = random.Random(seed)
    frames = []
    mx, my, pressed = 400.0, 300.0, False
    for _ in range(n):
        if rnd.random() < 0.15:
            pressed = not pressed
        if pressed:
            mx += rnd.uniform(-30, 30)
            my += rnd.uniform(-20, 20)
        keys = "".join(k for k in "WSADUCF" if rnd.random() < (0.05 if k == "F" else 0.2))
        focus = rnd.choice([0.0, 0.0, 0.0, 0.1, -0.1])
        frames.append((rnd.choice([1 / 60, 1 / 30, 0.05]), keys, round(mx, 2), round(my, 2), pressed, focus))
    return frames


def test_cpp_camera_controller_matches_python_controller(cli, crt, tmp_path):
    """CRT::Camera::updateCamera (C++, float, cosf/sinf like the reference's host code) against controller.py
    (numpy float32, cos/sin through double) on 400 scripted frames: same quality switching and motion flags
    exactly, poses within 1e-4 (the two differ only in the last ulp of the trigonometric functions, which the
    position integrates over the frames)."""
    ct = importlib.import_module("raytracer-cuda_b200.controller")
    frames = _script(400, 7)
    path = tmp_path / "script.txt"
    with open(path, "w") as f:
        for dt, keys, mx, my, pressed, focus in frames:
            f.write("%.9g %s %.9g %.9g %d %.9g\n" % (dt, keys or "-", mx, my, int(pressed), focus))
    out = subprocess.run([cli, "--controller-trace", str(path), "--width", "800", "--height", "600"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    rows = [line.split() for line in out.stdout.strip().split("\n")]
    assert len(rows) == len(frames)
    cam = crt.Camera()
    cam.position[:] = (0.0, 4.0, 4.0)
    cam.world_up[:] = (0.0, 1.0, 0.0)
    cam.yaw_deg, cam.pitch_deg, cam.vfov_deg, cam.aspect, cam.aperture = -90.0, 0.0, 90.0, 800 / 600, 0.0
    cam.focus_dist = float(np.sqrt(np.float32(32.0)))
    names = {"W": "W", "S": "S", "A": "A", "D": "D", "U": "Space", "C": "LControl", "F": "F"}
    c = ct.CameraController(cam, 800, 600)
    moved = 0
    for (dt, keys, mx, my, pressed, focus), row in zip(frames, rows):
        if focus:
            c.adjust_focus_distance(focus)
        spp = c.update(np.float32(dt), tuple(names[k] for k in keys), (mx, my), pressed)
        got = [float(x) for x in row[:6]]
        want = [cam.position[0], cam.position[1], cam.position[2], cam.yaw_deg, cam.pitch_deg, cam.focus_dist]
        assert np.allclose(got, want, rtol=0, atol=1e-4), (got, want)
        assert int(row[6]) == spp
        assert int(row[7]) == int(c.in_motion())
        moved += int(c.in_motion())
    assert 0 < moved < len(frames)


def _read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = [int(x) for x in f.readline().split()]
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data


@pytest.mark.gpu
def test_cli_still_frame_equals_binding(cli, crt, tmp_path):
    """One still frame through the C++ host (crt_cli -> CRT::CUDARenderer) and the same frame through the ctypes
    binding: the PPM and PFM files are byte-identical."""
    W = H = 96
    ppm_c, pfm_c = tmp_path / "c.ppm", tmp_path / "c.pfm"
    out = subprocess.run([cli, "--scene", "cornell", "--width", str(W), "--height", str(H), "--spp", "8", "--bounces", "8", "--mode", "wavefront",
                          "--ppm", str(ppm_c), "--pfm", str(pfm_c)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    info = json.loads(out.stdout.strip().split("\n")[-1])
    assert info["frames"] == 1 and info["rays"] > W * H * 8
    sc = crt.builtin_scene("cornell")
    r = crt.Renderer(W, H, mode=crt.MODE_WAVEFRONT)
    r.set_scene(sc)
    cam = sc.camera
    cam.aspect = W / H
    r.update_camera(cam)
    r.initialize(SEED, 0)
    st = r.render(8, 8)
    assert int(st.rays) == info["rays"]
    ppm_p, pfm_p = tmp_path / "p.ppm", tmp_path / "p.pfm"
    r.write_ppm(str(ppm_p))
    r.write_pfm(str(pfm_p))
    r.close()
    sc.free()
    assert open(ppm_c, "rb").read() == open(ppm_p, "rb").read()
    assert open(pfm_c, "rb").read() == open(pfm_p, "rb").read()
    assert np.isfinite(_read_pfm(pfm_c)).all()


@pytest.mark.gpu
def test_cli_progressive_frames(cli):
    """Raytracer::run without the window: 6 frames, the camera turning (1 spp per frame while it moves)."""
    out = subprocess.run([cli, "--scene", "cornell", "--width", "128", "--height", "72", "--frames", "6", "--orbit", "0.5", "--bounces", "8"],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    info = json.loads(out.stdout.strip().split("\n")[-1])
    assert info["frames"] == 6 and info["rays"] >= 6 * 128 * 72


@pytest.mark.gpu
def test_cli_two_gpus_equals_device_group(cli, crt, tmp_path):
    """crt_cli --gpus 2 (CRT::CUDARendererGroup -> crt_group_*): the frame file equals the one the ctypes DeviceGroup
    writes from GPU 0 after the same sharded render, with the peer-memory kernel and with ncclReduce."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    W, H = 160, 96
    sc = crt.builtin_scene("cornell")
    cam = sc.camera
    cam.aspect = W / H
    g = crt.DeviceGroup(W, H, [0, 1], mode=crt.MODE_WAVEFRONT)
    g.set_scene(sc)
    g.update_camera(cam)
    g.initialize(SEED)
    st = g.render(9, 8)
    ppm_p = tmp_path / "p.ppm"
    g.renderer(0).write_ppm(str(ppm_p))
    g.close()
    sc.free()
    for extra in ([], ["--nccl"]):
        ppm_c = tmp_path / ("c%d.ppm" % len(extra))
        out = subprocess.run([cli, "--scene", "cornell", "--width", str(W), "--height", str(H), "--spp", "9", "--bounces", "8", "--mode", "wavefront",
                              "--gpus", "2", "--ppm", str(ppm_c)] + extra, capture_output=True, text=True)
        assert out.returncode == 0, out.stderr
        info = json.loads(out.stdout.strip().split("\n")[-1])
        assert info["gpus"] == 2 and info["rays"] == int(st.rays)
        assert info["reduce"] == ("ncclReduce" if extra else "peer-memory kernel")
        assert open(ppm_c, "rb").read() == open(ppm_p, "rb").read()
