"""T7: the render-command wire protocol against elevenrender_b200/er_server — a Python client
that speaks what ElevenBlender speaks: 1024-byte NUL-padded JSON headers + payloads, CLI-string commands.

CPU (-m "not gpu"): framing, the accept-time STATUS ok, every load command answers STATUS ok, --get_sycl_info
returns the {"devices": [...]} schema, unknown commands are ignored, close_session ends the session.
GPU (-m gpu): config + camera + HDRI + materials + OBJ object -> --start -> poll --get_info up to N+1 ->
--get_pass beauty returns w*h*16 bytes equal to the direct C-ABI render of the same scene."""
import json
import os
import socket
import struct
import subprocess
import time

import numpy as np
import pytest

import common
from elevenrender_b200 import capi, host, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = 1024


class Client:
    def __init__(self, port):
        self.s = socket.create_connection(("127.0.0.1", port), timeout=60)

    def _read(self, n):
        buf = bytearray()
        while len(buf) < n:
            c = self.s.recv(n - len(buf))
            if not c:
                raise EOFError
            buf += c
        return bytes(buf)

    def recv(self):
        h = json.loads(self._read(HEADER).split(b"\0", 1)[0].decode())
        return h, (self._read(h["data_size"]) if h["data_size"] else b"")

    def send(self, typ, fmt, payload):
        if isinstance(payload, str):
            payload = payload.encode()
        h = json.dumps({"type": typ, "data_format": fmt, "data_size": len(payload)}).encode()
        self.s.sendall(h + b"\0" * (HEADER - len(h)) + payload)

    def command(self, text, *data):
        self.send("command", "string", text)
        for fmt, payload in data:
            self.send("data", fmt, payload)
        return self.recv()

    def close(self):
        self.send("status", "string", "close_session")
        self.s.close()


@pytest.fixture()
def server():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "elevenrender_b200", "csrc"), "server"])
    p = subprocess.Popen([os.path.join(ROOT, "elevenrender_b200", "er_server"), "--port", "0"], stdout=subprocess.PIPE,
                         stderr=subprocess.PIPE, text=True)
    port = int(p.stdout.readline().split()[1])
    yield port
    p.terminate()
    try:
        p.wait(timeout=10)
    except subprocess.TimeoutExpired:
        p.kill()


CAMERA = {"position": {"x": 0.0, "y": 0.0, "z": -1.0}, "rotation": {"x": 0.0, "y": 0.0, "z": 0.0}, "aperture": 2.8,
          "bokeh": False, "focus_distance": 4.0, "focal_length": 0.035, "sensor_width": 0.036, "sensor_height": 0.024}


def ok(reply):
    h, body = reply
    assert h == {"type": "status", "data_format": "string", "data_size": 2} and body == b"ok", (h, body)


def test_framing_and_load_commands(server):
    c = Client(server)
    ok(c.recv())  # the server greets with STATUS ok (src/main.cpp:211)
    ok(c.command("--load_config", ("json", json.dumps({"x_res": 32, "y_res": 16, "sample_target": 2, "denoise": False,
                                                        "device": "NVIDIA B200|NVIDIA CUDA", "block_size": 8}))))
    ok(c.command("--load_camera", ("json", json.dumps(CAMERA))))
    tex = np.linspace(0, 1, 4 * 2 * 3, dtype=np.float32)
    ok(c.command("--load_texture", ("json", json.dumps({"name": "t0", "width": 4, "height": 2, "channels": 3, "color_space": "LINEAR"})),
                 ("float3", tex.tobytes())))
    ok(c.command("--load_hdri --mirror_x", ("json", json.dumps({"name": "env", "width": 4, "height": 2, "channels": 3, "color_space": "LINEAR"})),
                 ("float3", tex.tobytes())))
    ok(c.command("--load_brdf_material", ("json", json.dumps({"name": "red", "albedo": {"r": 0.8, "g": 0.1, "b": 0.1}, "roughness": 0.4,
                                                                "albedo_map": "t0"}))))
    obj, mtl = scenes.cornell_obj_text()
    ok(c.command("--load_object --recompute_normals", ("string", obj), ("string", mtl)))
    # malformed JSON is logged and still answered with ok (src/CommandManager.cpp:121-128)
    ok(c.command("--load_camera", ("json", "{not json")))
    h, body = c.command("--get_sycl_info")
    assert h["type"] == "data" and h["data_format"] == "json" and h["data_size"] == len(body)
    info = json.loads(body)
    assert isinstance(info["devices"], list)
    for d in info["devices"]:
        assert set(d) == {"name", "platform", "memory", "max_compute_units", "is_compatible", "online_compiler", "type"}
    # an unknown option is not answered; the session goes on
    c.send("command", "string", "--no_such_option")
    h, body = c.command("--get_sycl_info")
    assert h["type"] == "data"
    c.close()
    # the server takes the next session
    c2 = Client(server)
    ok(c2.recv())
    c2.close()


def test_command_tokens_follow_std_quoted(server):
    """str_to_argv (src/main.cpp:13-24): std::quoted words; --get_pass takes its value as the next word or after '='"""
    c = Client(server)
    ok(c.recv())
    ok(c.command("--load_config", ("json", json.dumps({"x_res": 4, "y_res": 2, "sample_target": 1, "denoise": True, "device": "",
                                                        "block_size": 8}))))
    for text in ('--get_pass beauty', '--get_pass "beauty"', '--get_pass=normal'):
        h, body = c.command(text)
        assert h == {"type": "data", "data_format": "float4", "data_size": 4 * 2 * 16}
        a = np.frombuffer(body, dtype=np.float32).reshape(2, 4, 4)
        assert np.all(a[..., :3] == 0) and np.all(a[..., 3] == 1)  # nothing rendered: zeros, alpha forced by the denoise flag
    c.close()


@pytest.mark.gpu
def test_render_over_the_protocol(gpu, server):
    spec = common.make_spec("cornell")
    w, h, spp = 64, 48, 3
    obj, mtl = scenes.cornell_obj_text()
    c = Client(server)
    ok(c.recv())
    ok(c.command("--load_config", ("json", json.dumps({"x_res": w, "y_res": h, "sample_target": spp, "denoise": False,
                                                        "device": f"{gpu['name']}|{gpu['platform']}", "block_size": 8}))))
    cam = dict(CAMERA)
    cam["position"] = {"x": 0.0, "y": 1.0, "z": -3.4}
    ok(c.command("--load_camera", ("json", json.dumps(cam))))
    env = (np.linspace(0.1, 1.0, 16 * 8 * 3, dtype=np.float32) ** 2).astype(np.float32)
    ok(c.command("--load_hdri", ("json", json.dumps({"name": "env", "width": 16, "height": 8, "channels": 3, "color_space": "LINEAR"})),
                 ("float3", env.tobytes())))
    mats = [{"name": "red", "albedo": {"r": 0.63, "g": 0.065, "b": 0.05}}, {"name": "green", "albedo": {"r": 0.14, "g": 0.45, "b": 0.091}},
            {"name": "light", "albedo": {"r": 0.78, "g": 0.78, "b": 0.78}, "emission": {"r": 8.0, "g": 8.0, "b": 8.0}}]
    for m in mats:
        ok(c.command("--load_brdf_material", ("json", json.dumps(m))))
    ok(c.command("--load_object", ("string", obj), ("string", mtl)))
    ok(c.command("--start"))
    t0 = time.time()
    while True:
        hd, body = c.command("--get_info")
        assert hd["data_format"] == "json"
        n = json.loads(body)["samples"]
        if n >= spp + 1:  # dev_samples starts at 1
            break
        assert time.time() - t0 < 60
        time.sleep(0.01)
    assert n == spp + 1
    hd, body = c.command("--get_pass beauty")
    assert hd == {"type": "data", "data_format": "float4", "data_size": w * h * 16}
    got = np.frombuffer(body, dtype=np.float32).reshape(h, w, 4)
    hd, body = c.command("--get_pass normal")
    got_n = np.frombuffer(body, dtype=np.float32).reshape(h, w, 4)
    c.close()

    # the same scene through the Python mirror of the C ABI: our loader's triangles, the same pairing by name
    tris, names = capi.load_obj(obj, mtl)
    table = {"default": 0, "red": 1, "green": 2, "light": 3}
    tris[:, 38] = np.array([table.get(nm, 0) for nm in names], dtype=np.int32).view(np.float32)
    sc = host.Scene(w, h)
    for m in mats:
        kw = {"albedo": tuple(m["albedo"][k] for k in "rgb")}
        if "emission" in m:
            kw["emission"] = tuple(m["emission"][k] for k in "rgb")
        # parse_materialjson leaves ax = ay = max(0.001, roughness / sqrt(1 - 0.9 anisotropic)) = 1 (src/CommandManager.cpp:108-110)
        sc.addMaterial(host.material_words(ax=1.0, ay=1.0, **kw), m["name"])
    sc.addMeshObject(tris)
    shifted = np.roll(env.reshape(8, 16, 3), 8, axis=1)  # pixel_shift(0.5, 0) (src/CommandManager.cpp:187)
    sc.addHDRI(host.HDRI(host.Texture("env", 16, 8, 3, shifted.reshape(-1))))
    sc.camera = host.camera_words(position=(0.0, 1.0, -3.4), rotation=(0.0, 0.0, 0.0), aperture=2.8, bokeh=False, focusDistance=4.0,
                                  focalLength=0.035, sensorWidth=0.036, sensorHeight=0.024)
    rm = host.RenderingManager(host.RenderParameters(w, h, spp, rng_order=host.ER_RNG_LTR))
    rm.upload(sc)
    rm.render_samples(spp)
    want, want_n = rm.get_pass("beauty"), rm.get_pass("normal")
    rm.close()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(got_n.view(np.uint32), want_n.view(np.uint32))
    assert got[..., :3].max() > 0
