"""CPU: the C++ host (host/*.cpp: Conf / .ini / command-line precedence, the .mat text parser, steering, the true
vehicle, sensor, noise draws and the known data association) against the UNMODIFIED reference's own functions
(oracle/_ref/libslamref.so through oracle/ref_shim_host.cpp).  Both sides expose the same C signatures.  Skipped when
the reference library is absent (the GPU box); there tests/test_gpu_host_app.py holds the whole program against the
reference program's golden traces."""
import ctypes as C
import glob
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libslamref.so")
REF_DATA = "/root/reference/data"
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def _declare(L, pre):
    g = lambda n: getattr(L, pre + n)  # noqa: E731
    g("conf_dump").argtypes = [C.c_char_p, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.c_int, C.c_char_p, C.c_int]
    g("read_map").argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_void_p, C.c_char_p,
                              C.c_int]
    g("add_control_noise").argtypes = [C.c_float, C.c_float, f32p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    g("predict_true_position").argtypes = [f32p, C.c_float, C.c_float, C.c_float, C.c_float]
    g("set_map").argtypes = [f32p, C.c_int, f32p, C.c_int]
    g("update_steering").argtypes = [f32p, C.POINTER(C.c_int), C.c_float, C.POINTER(C.c_float), C.c_float, C.c_float,
                                     C.c_float]
    g("get_observations").argtypes = [f32p, C.c_float, C.c_int, f32p, i32p, f32p]
    g("data_association_known").argtypes = [f32p, i32p, C.c_int, i32p, C.c_int, C.c_int, f32p, i32p, f32p,
                                            C.POINTER(C.c_int)]
    return g


@pytest.fixture(scope="module")
def sides():
    if not os.path.exists(REF_LIB):
        pytest.skip("oracle/_ref/libslamref.so not built here")
    import slam_b200
    slam_b200.build()
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host"), "libslamhost.so"])
    H = C.CDLL(os.path.join(ROOT, "host", "libslamhost.so"))
    R = C.CDLL(REF_LIB)
    R.slamref_rand_set_stream.argtypes = [i32p, C.c_long]
    return _declare(H, "slamhost_"), _declare(R, "slamref_"), H, R


def conf_dump(g, ini, over):
    keys = (C.c_char_p * max(1, len(over)))(*[k.encode() for k in over])
    vals = (C.c_char_p * max(1, len(over)))(*[str(v).encode() for v in over.values()])
    out = C.create_string_buffer(4096)
    g("conf_dump")(ini.encode(), keys, vals, len(over), out, 4096)
    return out.value.decode()


def read_map(g, path):
    nl, nw = C.c_int(), C.c_int()
    err = C.create_string_buffer(512)
    assert g("read_map")(path.encode(), C.byref(nl), C.byref(nw), None, None, err, 512) == 0, err.value
    lm, wp = np.zeros((nl.value, 2), np.float32), np.zeros((nw.value, 2), np.float32)
    g("read_map")(path.encode(), C.byref(nl), C.byref(nw), lm.ctypes.data, wp.ctypes.data, err, 512)
    return lm, wp


def test_conf_defaults_ini_and_command_line_precedence(sides):
    h, r, _, _ = sides
    inis = sorted(glob.glob(os.path.join(REF_DATA, "*.ini"))) + [os.path.join(HERE, "data", "ring_small.ini"), ""]
    assert len(inis) >= 2
    overrides = [{}, {"NPARTICLES": 5000}, {"Vtrue": "1.25", "SWITCH_SEED_RANDOM": 7, "sigmaT": "0.05", "method": "EKF1"},
                 {"NEFFECTIVE": 10, "MAX_RANGE": "17.5", "unknown_key": "x"}]
    for ini in inis:
        for over in overrides:
            a, b = conf_dump(h, ini, over), conf_dump(r, ini, over)
            assert a == b, (ini, over, a, b)
    # NEFFECTIVE is NOT recomputed from an overridden NPARTICLES (core.cpp:1011-1012)
    assert "NPARTICLES=5000 NEFFECTIVE=75" in conf_dump(h, "", {"NPARTICLES": 5000})


def test_map_files(sides, tmp_path):
    h, r, _, _ = sides
    maps = sorted(glob.glob(os.path.join(REF_DATA, "*.mat"))) + [os.path.join(HERE, "data", "ring_small.mat")]
    odd = tmp_path / "odd.mat"  # comments, blank lines, extra tokens on a line, tabs
    # (lm first: the reference reads the wp rows with the LANDMARK row count, core.cpp:948, so a wp block in front of
    # the lm block yields uninitialised waypoints there; host/conf.cpp refuses such a file instead)
    odd.write_text("# c\n\nlm 2 3\n0 0\n  1 1  \n-5.25 1e-3 9 9\n# c\nwp 2 2\n1.5\t-2 ignored\n3e1 4\n")
    maps.append(str(odd))
    for m in maps:
        (la, wa), (lb, wb) = read_map(h, m), read_map(r, m)
        assert la.shape == lb.shape and wa.shape == wb.shape and len(la) > 0 and len(wa) > 0, m
        assert np.array_equal(la, lb) and np.array_equal(wa, wb), m
    nl, nw = C.c_int(), C.c_int()
    err = C.create_string_buffer(512)
    bad = tmp_path / "bad.mat"
    bad.write_text("lm 2 2\n1 2\n")
    assert h("read_map")(str(bad).encode(), C.byref(nl), C.byref(nw), None, None, err, 512) == -1 and err.value
    assert h("read_map")(b"/nonexistent.mat", C.byref(nl), C.byref(nw), None, None, err, 512) == -1


def test_drive_steering_sensor_noise_and_association_bitwise(sides):
    """A 600-tick drive through both sides with ONE rand() stream: steering angle, true pose, noisy controls, the set of
    visible landmarks, noisy range / bearing and the known association must agree bit for bit."""
    h, r, H, R = sides
    lm, wp = read_map(h, os.path.join(HERE, "data", "ring_small.mat"))
    for g in (h, r):
        g("set_map")(lm, len(lm), wp, len(wp))
    libc = C.CDLL(None)
    libc.srand(99)
    stream = np.array([libc.rand() for _ in range(60000)], np.int32)
    R.slamref_rand_set_stream(stream, len(stream))
    H.slamhost_srand(99)
    Q = np.array([0.04, 0, 0, 0.0012], np.float32)
    Rn = np.array([0.01, 0, 0, 0.0003], np.float32)
    st = []
    for g in (h, r):
        st.append(dict(x=np.zeros(3, np.float32), G=C.c_float(0), iwp=C.c_int(0), table=-np.ones(len(lm), np.int32),
                       Nf=0))
    n_seen = 0
    for tick in range(600):
        outs = []
        for g, s in zip((h, r), st):
            g("update_steering")(s["x"], C.byref(s["iwp"]), 1.0, C.byref(s["G"]), 0.5, 0.6, 0.025)
            g("predict_true_position")(s["x"], 2.0, s["G"].value, 2.0, 0.025)
            vn, gn = C.c_float(), C.c_float()
            g("add_control_noise")(2.0, s["G"].value, Q, C.byref(vn), C.byref(gn))
            o = [s["G"].value, s["iwp"].value, s["x"].copy(), vn.value, gn.value]
            if tick % 8 == 7:
                ids, z = np.zeros(len(lm), np.int32), np.zeros((len(lm), 2), np.float32)
                k = g("get_observations")(s["x"], 12.0, 1, Rn, ids, z)
                zf, idf, zn = np.zeros((k + 1, 2), np.float32), np.zeros(k + 1, np.int32), np.zeros((k + 1, 2), np.float32)
                kn = C.c_int()
                K = g("data_association_known")(z[:k].copy(), ids[:k].copy(), k, s["table"], len(lm), s["Nf"], zf, idf, zn,
                                                C.byref(kn))
                s["Nf"] += kn.value
                o += [k, ids[:k].copy(), z[:k].copy(), K, kn.value, zf[:K].copy(), idf[:K].copy(), zn[:kn.value].copy(),
                      s["table"].copy()]
                n_seen = s["Nf"]
            outs.append(o)
        for a, b in zip(*outs):
            assert np.array_equal(np.asarray(a), np.asarray(b)), tick
    assert n_seen > 10


def test_config_fixture_carries_the_reference_inputs(sides, tmp_path):
    """tests/golden/config_traces.npz rebuilds BASELINE configs[0..2]' input files on the GPU box: the rebuilt .mat /
    .ini must parse (by the reference's own parser and by the host's) to what the reference's files parse to, apart
    from the recorded fixed seed."""
    from test_gpu_configs import write_inputs
    h, r, _, _ = sides
    g = np.load(os.path.join(HERE, "golden", "config_traces.npz"))
    for tag, name in (("cfg0", "example_webmap"), ("cfg1", "example_loop1"), ("cfg2", "example_loop902")):
        d = tmp_path / tag
        d.mkdir()
        write_inputs(g, tag, str(d))
        la, wa = read_map(r, os.path.join(REF_DATA, name + ".mat"))
        for side in (h, r):
            lb, wb = read_map(side, str(d / "map.mat"))
            assert np.array_equal(la, lb) and np.array_equal(wa, wb), (tag, name)
        seed = [s for s in g[tag + "_ini"] if str(s).startswith("SWITCH_SEED_RANDOM")][0].split("=")[1].strip()
        want = conf_dump(r, os.path.join(REF_DATA, name + ".ini"), {"SWITCH_SEED_RANDOM": seed})
        assert conf_dump(h, str(d / "map.ini"), {}) == want and conf_dump(r, str(d / "map.ini"), {}) == want
