"""CPU: the GUI feed (host/zmq_plot.cpp) on the wire.  A pyzmq PAIR socket BOUND where the reference's slam-gui
listens stands in for the GUI; the host emits a scripted sequence through its zmq PlotSink and every multipart message
must equal, frame for frame, what the reference's NetworkPlot (plotting/NetworkPlot.cpp:22-218) would put on the
wire through zmqpp (libs/zmqpp/message.cpp:225-328: one frame per value, 32-bit integers in network order, float /
double big-endian, strings raw, the method name first)."""
import ctypes as C
import glob
import os
import struct
import subprocess
import threading

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def i32(v):
    return struct.pack(">i", v)


def u32(v):
    return struct.pack(">I", v)


def f32(v):
    return struct.pack(">f", v)


def f64(v):
    return struct.pack(">d", v)


def xy_msg(name, xs, ys):
    return [name.encode(), i32(len(xs))] + [f64(v) for v in xs] + [i32(len(ys))] + [f64(v) for v in ys]


def mat_msg(name, M, tail=()):
    M = np.asarray(M, np.float32)
    return [name.encode(), u32(M.shape[0]), u32(M.shape[1])] + [f32(float(v)) for v in M.ravel()] + list(tail)


def test_wire_frames_match_the_reference_protocol():
    zmq = pytest.importorskip("zmq")
    libs = glob.glob(os.path.join(os.path.dirname(os.path.dirname(zmq.__file__)), "pyzmq.libs", "libzmq-*.so*"))
    if not libs and not any(os.path.exists(p) for p in ("/usr/lib/x86_64-linux-gnu/libzmq.so.5",)):
        pytest.skip("no libzmq on this box")
    if libs:
        os.environ["SLAMGPU_LIBZMQ"] = libs[0]
    import slam_b200
    slam_b200.build()
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host"), "libslamhost.so"])
    H = C.CDLL(os.path.join(ROOT, "host", "libslamhost.so"))
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    H.slamhost_set_map.argtypes = [f32p, C.c_int, f32p, C.c_int]
    H.slamhost_zmq_selftest.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    lm = np.array([[-3, 4], [10, -2], [6, 8]], np.float32)
    wp = np.array([[0, 0], [5, 5]], np.float32)
    H.slamhost_set_map(lm, 3, wp, 2)

    ctx = zmq.Context()
    gui = ctx.socket(zmq.PAIR)
    port = gui.bind_to_random_port("tcp://127.0.0.1")
    got = []

    def serve():
        while True:
            if not gui.poll(5000):
                return
            m = gui.recv_multipart()
            got.append(m)
            if m[0] == b"endPlot":
                return

    t = threading.Thread(target=serve)
    t.start()
    err = C.create_string_buffer(256)
    assert H.slamhost_zmq_selftest(("tcp://127.0.0.1:%d" % port).encode(), err, 256) == 0, err.value
    t.join(20)
    gui.close(0)
    ctx.term()

    # what the reference would have sent: app ctor + configurePlot, then one FastSLAM2 iteration, then endPlot
    xs, ys = [-3.0, 10.0, 6.0, 0.0, 5.0], [4.0, -2.0, 8.0, 0.0, 5.0]
    xmin, xmax, ymin, ymax = min(xs), max(xs), min(ys), max(ys)
    xt, xe = (1.0, 2.0, 0.5), (1.25, 1.75, 0.45)
    c, s = np.cos(np.float32(0.5)), np.sin(np.float32(0.5))
    z = np.array([[5.0, 0.25], [7.5, -1.0]], np.float32)
    gx, gy = z[:, 0] * np.cos(z[:, 1]), z[:, 0] * np.sin(z[:, 1])
    lines = np.stack([np.full(2, 1.0, np.float32), np.full(2, 2.0, np.float32), (c * gx - s * gy) + np.float32(1.0),
                      (s * gx + c * gy) + np.float32(2.0)]).astype(np.float32)
    px, py = [np.float32(v) for v in (0.1, 0.2, 0.3)], [np.float32(v) for v in (1.1, 1.2, 1.3)]
    fx, fy = [10, 20, 12, 22, 14, 24], [11, 21, 13, 23, 15, 25]   # particle-major, landmark-minor
    want = [
        [b"setSimulationName", b"selftest"],
        [b"setCarSize", f64(4.0), u32(0)],
        [b"setCarSize", f64(4.0), u32(1)],
        xy_msg("setWaypoints", [0.0, 5.0], [0.0, 5.0]),
        xy_msg("setLandmarks", [-3.0, 10.0, 6.0], [4.0, -2.0, 8.0]),
        [b"setPlotRange", f64(xmin - (xmax - xmin) * 0.05), f64(xmax + (xmax - xmin) * 0.05),
         f64(ymin - (ymax - ymin) * 0.05), f64(ymax + (ymax - ymin) * 0.05)],
        [b"addTruePosition", f64(0), f64(0)],
        [b"setCarTruePosition", f64(0), f64(0), f64(0)],
        [b"addEstimatedPosition", f64(0), f64(0)],
        [b"setCarEstimatedPosition", f64(0), f64(0), f64(0)],
        [b"plot"],
        None,  # setCovEllipse: checked structurally below
        [b"loopTime", u32(1234)],
        xy_msg("setParticles", [float(v) for v in px], [float(v) for v in py]),
        xy_msg("setFeatureParticles", [float(v) for v in fx], [float(v) for v in fy]),
        [b"addTruePosition", f64(xt[0]), f64(xt[1])],
        [b"addEstimatedPosition", f64(xe[0]), f64(xe[1])],
        [b"setCarTruePosition", f64(xt[0]), f64(xt[1]), f64(float(np.float32(xt[2])))],
        [b"setCarEstimatedPosition", f64(xe[0]), f64(xe[1]), f64(xe[2])],
        None,  # setLaserLines: compared numerically (libm cosf / sinf vs numpy may differ by an ulp)
        [b"plot"],
        [b"endPlot"],
    ]
    assert [m[0] for m in got] == [w[0] if w else n for w, n in zip(want, [None] * 11 + [b"setCovEllipse"] + [None] * 7 +
                                                                    [b"setLaserLines", None, None])]
    for g, w in zip(got, want):
        if w is not None:
            assert g == w, (g[0], g, w)
    ell = got[11]
    assert ell[1:3] == [u32(2), u32(17)] and len(ell) == 3 + 34 + 1 and ell[-1] == i32(0)
    E = np.array([struct.unpack(">f", b)[0] for b in ell[3:-1]]).reshape(2, 17)
    assert np.allclose(E[:, 0], E[:, 16], atol=1e-5)                           # closed curve
    d = E - np.array([[1.0], [2.0]])
    Pinv = np.linalg.inv(np.array([[0.04, 0.01], [0.01, 0.09]]))
    assert np.allclose(np.einsum("ik,ij,jk->k", d, Pinv, d), 4.0, rtol=1e-3)   # the 2-sigma contour of P
    las = got[19]
    assert las[1:3] == [u32(4), u32(2)] and len(las) == 3 + 8
    L = np.array([struct.unpack(">f", b)[0] for b in las[3:]], np.float32).reshape(4, 2)
    assert np.allclose(L, lines, rtol=0, atol=2e-6)
