"""The offline bounce driver `za_bounce` (host/za_bounce.cpp + host/za_wav.h): WAV
decode / 24-bit encode on the CPU, and -- on the GPU -- a whole bounce of ragged WAV files against the reference's
compiled DSP rendered over the same samples."""
import json
import struct
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parent.parent
PKG = REPO / "zorakaudio-experimental-plugins_b200"


@pytest.fixture(scope="module")
def bounce(built_lib):
    sys.path.insert(0, str(PKG))
    from za_gpu import build
    return str(build.build_bounce())


def write_wav(path, x, sr, kind):
    """x [ch][frames] float64 in [-1, 1).  kind: i16 | i24 | i32 | f32 | f64 | ext24 (WAVE_FORMAT_EXTENSIBLE).
    Returns the float32 samples a correct reader must decode."""
    ch, n = x.shape
    if kind in ("i16", "i24", "ext24", "i32"):
        bits = {"i16": 16, "i24": 24, "ext24": 24, "i32": 32}[kind]
        q = np.clip(np.round(x * (1 << (bits - 1))), -(1 << (bits - 1)), (1 << (bits - 1)) - 1).astype(np.int64)
        inter = q.T.reshape(-1)
        if bits == 16:
            raw = inter.astype("<i2").tobytes()
        elif bits == 32:
            raw = inter.astype("<i4").tobytes()
        else:
            b = inter.astype("<i4").view(np.uint8).reshape(-1, 4)[:, :3]
            raw = b.tobytes()
        want = (q.astype(np.float64) / float(1 << (bits - 1))).astype(np.float32) if bits < 32 else (q.astype(np.float64) / 2147483648.0).astype(np.float32)
        tag = 1
    else:
        bits = 32 if kind == "f32" else 64
        arr = x.T.reshape(-1).astype("<f4" if bits == 32 else "<f8")
        raw = arr.tobytes()
        want = x.astype(np.float32) if bits == 32 else x.astype(np.float64).astype(np.float32)
        tag = 3
    ba = ch * bits // 8
    if kind == "ext24":
        guid = struct.pack("<H", 1) + bytes.fromhex("000000001000800000aa00389b71")
        fmt = struct.pack("<HHIIHHHHI", 0xFFFE, ch, sr, sr * ba, ba, bits, 22, bits, 0) + guid
    else:
        fmt = struct.pack("<HHIIHH", tag, ch, sr, sr * ba, ba, bits)
    junk = b"LIST" + struct.pack("<I", 5) + b"hello" + b"\0"            # odd-sized chunk before fmt: padding must be skipped
    body = b"WAVE" + junk + b"fmt " + struct.pack("<I", len(fmt)) + fmt + b"data" + struct.pack("<I", len(raw)) + raw + (b"\0" if len(raw) & 1 else b"")
    Path(path).write_bytes(b"RIFF" + struct.pack("<I", len(body)) + body)
    return want


def read_wav24(path):
    b = Path(path).read_bytes()
    assert b[:4] == b"RIFF" and b[8:16] == b"WAVEfmt "
    tag, ch, sr, _, ba, bits = struct.unpack("<HHIIHH", b[20:36])
    assert (tag, bits, ba) == (1, 24, 3 * ch)
    assert b[36:40] == b"data"
    n = struct.unpack("<I", b[40:44])[0]
    raw = np.frombuffer(b[44:44 + n], np.uint8).reshape(-1, 3).astype(np.int32)
    v = raw[:, 0] | (raw[:, 1] << 8) | (raw[:, 2] << 16)
    v = np.where(v >= 1 << 23, v - (1 << 24), v)
    return v.reshape(-1, ch).T, sr


def fnv1a(f32):
    h = 1469598103934665603
    for byte in np.ascontiguousarray(f32, dtype="<f4").tobytes():
        h = ((h ^ byte) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


@pytest.mark.parametrize("kind", ["i16", "i24", "i32", "f32", "f64", "ext24"])
def test_wav_reader_decodes_every_format(bounce, tmp_path, kind):
    rng = np.random.default_rng(5)
    x = rng.uniform(-0.99, 0.99, (3, 1001))
    want = write_wav(tmp_path / "a.wav", x, 44100, kind)
    r = subprocess.run([bounce, "--wav-info", str(tmp_path / "a.wav")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout)
    assert (info["channels"], info["frames"], info["sample_rate"]) == (3, 1001, 44100.0)
    assert info["float"] == (kind in ("f32", "f64"))
    assert info["fnv1a"] == fnv1a(want)


def test_cli_errors_are_loud(bounce, tmp_path):
    assert subprocess.run([bounce], capture_output=True).returncode == 2
    (tmp_path / "bad.wav").write_bytes(b"not a wave file at all")
    r = subprocess.run([bounce, "--wav-info", str(tmp_path / "bad.wav")], capture_output=True, text=True)
    assert r.returncode == 1 and "RIFF" in r.stderr
    write_wav(tmp_path / "a.wav", np.zeros((2, 64)), 48000, "i16")
    r = subprocess.run([bounce, "--plugin", "NoSuchPlugin", str(tmp_path / "a.wav")], capture_output=True, text=True)
    assert r.returncode == 1 and "NoSuchPlugin" in r.stderr
    r = subprocess.run([bounce, "--plugin", "EasyExpander", "--set", "99=1", str(tmp_path / "a.wav")], capture_output=True, text=True)
    assert r.returncode == 1 and "slider99" in r.stderr


def test_list_names_every_plugin(bounce):
    r = subprocess.run([bounce, "--list"], capture_output=True, text=True)
    assert r.returncode == 0
    for key in ("cascade8", "EasyExpander", "stft_mask2048", "RTT", "DDT"):
        assert f"\n{key}: " in "\n" + r.stdout


def test_preset_and_plugin_json(bounce, tmp_path):
    """APVTS-style state (<PARAM id="sliderN" value=...>, the reference's getStateInformation XML) and plugin.json slug:
    parsed on the host before any GPU work; --save-preset writes back the values the bounce uses."""
    write_wav(tmp_path / "a.wav", np.zeros((2, 64)), 48000, "i16")
    meta = {"name": "EasyExpander", "slug": "EasyExpander", "pluginCode": "EEXP", "pluginType": "jsfx", "entry": "src/EasyExpander.jsfx",
            "clapFeatures": ["audio-effect"]}
    (tmp_path / "plugin.json").write_text(json.dumps(meta, indent=2))
    (tmp_path / "state.xml").write_text('<?xml version="1.0"?><PARAMETERS><PARAM id="slider1" value="-41.5"/>'
                                        '<PARAM id="slider2" value="24"/><FILES f0="x.wav"/></PARAMETERS>')
    r = subprocess.run([bounce, "--plugin-json", str(tmp_path / "plugin.json"), "--preset", str(tmp_path / "state.xml"), "--set", "2=30",
                        "--save-preset", str(tmp_path / "used.xml"), "--out-dir", str(tmp_path / "o"), str(tmp_path / "a.wav")],
                       capture_output=True, text=True)
    used = (tmp_path / "used.xml").read_text()
    assert 'plugin="EasyExpander"' in used
    assert '<PARAM id="slider1" value="-41.5"/>' in used and '<PARAM id="slider2" value="30"/>' in used      # --set wins over the preset
    assert r.returncode in (0, 1)                        # 1 here = no CUDA device (see test_no_cpu_fallback)
    (tmp_path / "empty.xml").write_text("<PARAMETERS/>")
    r = subprocess.run([bounce, "--plugin", "EasyExpander", "--preset", str(tmp_path / "empty.xml"), str(tmp_path / "a.wav")], capture_output=True, text=True)
    assert r.returncode == 1 and "PARAM" in r.stderr


def test_plugin_json_is_validated_like_pluginlib(bounce, tmp_path):
    """The fields scripts/pluginlib.py:105-212 requires of a plugin.json, refused with its messages before any GPU work."""
    write_wav(tmp_path / "a.wav", np.zeros((2, 64)), 48000, "i16")
    good = {"name": "EasyExpander", "slug": "EasyExpander", "pluginCode": "EEXP", "pluginType": "jsfx", "entry": "src/EasyExpander.jsfx"}
    cases = [({"slug": ""}, "Missing 'slug'"), ({"pluginCode": "EEXPX"}, "exactly 4 characters"), ({"manufacturerCode": "Zo"}, "manufacturerCode"),
             ({"installPath": "x"}, "installPath"), ({"pluginType": "vst"}, "Invalid pluginType"), ({"pluginType": "faust"}, ".dsp"),
             ({"pluginType": "faust", "entry": "src/EasyExpander.dsp"}, "does not match")]
    for patch, msg in cases:
        (tmp_path / "plugin.json").write_text(json.dumps({**good, **patch}))
        r = subprocess.run([bounce, "--plugin-json", str(tmp_path / "plugin.json"), str(tmp_path / "a.wav")], capture_output=True, text=True)
        assert r.returncode == 1 and msg in r.stderr, (patch, r.stderr)
    r = subprocess.run([bounce, "--plugin", "EasyExpander", "--devices", "0-x", str(tmp_path / "a.wav")], capture_output=True, text=True)
    assert r.returncode == 2


def test_no_cpu_fallback(bounce, tmp_path):
    """Without a CUDA device a bounce fails with the library's message instead of rendering on the host."""
    import ctypes
    lib = ctypes.CDLL(str(PKG / "libza_gpu.so"))
    if lib.za_gpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    write_wav(tmp_path / "a.wav", np.zeros((2, 64)), 48000, "i16")
    r = subprocess.run([bounce, "--plugin", "EasyExpander", "--out-dir", str(tmp_path / "o"), str(tmp_path / "a.wav")], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.strip()


@pytest.mark.gpu
@pytest.mark.parametrize("key,precision", [("EasyExpander", "f64"), ("DDT", "f64"), ("EasyExpander", "f32")])
def test_bounce_matches_reference_render(bounce, tmp_path, key, precision):
    import za_oracle as zo
    from za_gpu import workloads as wl
    ref = zo.RefPlugin(key)
    sr = 48000
    lens = [30011, 48000, 17777]
    kinds = ["f32", "i16", "i24"]
    chans = [2, 2, 1]                                    # the mono file leaves the second input silent
    src = wl.programme_bursts(3, max(lens), n_ch=2).astype(np.float64)
    x = np.zeros((3, ref.n_in, max(lens)), np.float32)
    files = []
    for s in range(3):
        want = write_wav(tmp_path / f"in{s}.wav", src[s, :chans[s], :lens[s]], sr, kinds[s])
        x[s, :chans[s], :lens[s]] = want
        files.append(str(tmp_path / f"in{s}.wav"))
    p = np.tile(np.array([sl["default"] for sl in ref.sliders], np.float32), (3, 1))
    extra = []
    if key == "EasyExpander":
        p[:, 0] = -35.0
        if precision == "f64":
            extra = ["--set", "1=-35"]
        else:
            (tmp_path / "state.xml").write_text('<PARAMETERS><PARAM id="slider1" value="-35.0"/></PARAMETERS>')
            extra = ["--preset", str(tmp_path / "state.xml")]
    y, _ = ref.render(p, x, srate=float(sr), block=512)
    out = tmp_path / "out"
    r = subprocess.run([bounce, "--plugin", key, "--precision", precision, "--chunk-seconds", "0.25", "--out-dir", str(out), *extra, *files],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rep = json.loads((out / "report.json").read_text())
    assert rep["plugin"] == key and rep["streams"] == 3 and rep["kernelLaunches"] > 0 and rep["sampleRate"] == sr
    for s in range(3):
        got, gsr = read_wav24(rep["files"][s]["out"])
        assert gsr == sr and got.shape == (ref.n_out, lens[s])
        want = np.rint(np.clip(y[s, :, :lens[s]].astype(np.float64), -1.0, 1.0) * 8388607.0).astype(np.int64)
        d = np.abs(got - want)
        if precision == "f64":
            assert d.max() <= 1 and np.mean(d == 0) > 0.9999, (d.max(), np.mean(d == 0))
        else:                                           # fp32 mode: below -100 dBFS residual, i.e. well under one 24-bit step rms
            assert np.sqrt(np.mean((d / 8388607.0) ** 2)) < 1e-5


@pytest.mark.gpu
def test_bounce_resume_from_checkpoint_and_device_list(bounce, tmp_path):
    """A bounce interrupted after its first part (--state-out) and resumed by a second run (--state-in) gives the samples of the
    uninterrupted bounce; `--devices 0,0` shards the jobs over two handles (two GPUs where present) with identical results."""
    from za_gpu import workloads as wl
    sr, n1, n2 = 48000, 512 * 40, 512 * 33 + 100
    src = wl.programme_bursts(3, n1 + n2, n_ch=2).astype(np.float64)
    whole, first, second = [], [], []
    for s in range(3):
        write_wav(tmp_path / f"w{s}.wav", src[s], sr, "f32"); whole.append(str(tmp_path / f"w{s}.wav"))
        write_wav(tmp_path / f"a{s}.wav", src[s, :, :n1], sr, "f32"); first.append(str(tmp_path / f"a{s}.wav"))
        write_wav(tmp_path / f"b{s}.wav", src[s, :, n1:], sr, "f32"); second.append(str(tmp_path / f"b{s}.wav"))
    run = lambda args: subprocess.run([bounce, "--plugin", "SOMA", "--precision", "f64", "--chunk-seconds", "0.25"] + args, capture_output=True, text=True)
    r = run(["--out-dir", str(tmp_path / "ow")] + whole)
    assert r.returncode == 0, r.stderr
    r = run(["--out-dir", str(tmp_path / "oa"), "--state-out", str(tmp_path / "ck.bin"), "--devices", "0,0"] + first)
    assert r.returncode == 0, r.stderr
    r = run(["--out-dir", str(tmp_path / "ob"), "--state-in", str(tmp_path / "ck.bin"), "--devices", "0,0"] + second)
    assert r.returncode == 0, r.stderr
    for s in range(3):
        w, _ = read_wav24(tmp_path / "ow" / f"w{s}.{s}.wav")
        a, _ = read_wav24(tmp_path / "oa" / f"a{s}.{s}.wav")
        b, _ = read_wav24(tmp_path / "ob" / f"b{s}.{s}.wav")
        assert np.array_equal(np.concatenate([a, b], axis=1), w), s
    rep = json.loads((tmp_path / "oa" / "report.json").read_text())
    assert rep["devices"] == 2 and rep["oversampling"] == 1
