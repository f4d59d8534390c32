"""slamgpu_estimate_noalloc (include/slamgpu.h): the pose estimate with its scratch in a static __device__ array, for
callers that may not call cudaMalloc.  No counterpart in the reference,
so the oracle here is the library's own handle-based slamgpu_estimate on the same particles.

Collected last on purpose (file name): a failure here hides nothing.  Runs alone in
a fresh process and after the other GPU tests alike: it depends on no state they leave behind."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12  # the means are double sums of the same floats in another order: ~1e-16 relative per addition


@pytest.fixture(scope="module")
def sb():
    import slam_b200
    slam_b200.build()
    return slam_b200


def _particles(P, seed):
    rng = np.random.default_rng(seed)
    xv = np.stack([rng.normal(30.0, 5.0, P), rng.normal(-12.0, 3.0, P), rng.uniform(-3.1, 3.1, P)]).astype(np.float32)
    w = rng.random(P).astype(np.float32)
    # several particles share the maximum weight: the FIRST one's heading is the estimate
    top = np.float32(2.0)
    first = P // 3
    w[first] = top
    w[min(P - 1, first + 4097)] = top
    w[P - 1] = top
    return xv, w


@pytest.mark.parametrize("P", [100, 4096 * 3 + 17, 1 << 20])
def test_noalloc_estimate_matches_handle_estimate_twice(sb, P):
    import torch
    torch.cuda.set_device(0)
    f = sb.FastSLAM(sb.FASTSLAM2, P, 1, scan_mode=sb.SCAN_BLOCKED, spare_rows=1, log_epochs=4)
    out = torch.full((4,), float("nan"), dtype=torch.float64, device="cuda")
    for call in range(2):  # two calls in a row: the second finds the first one's partials and ticket in the scratch
        xv, w = _particles(P, seed=100 + call)
        f.upload(xv=xv, w=w)
        ex, ey, et = f.estimate()
        d_xv, d_w = torch.from_numpy(xv).cuda(), torch.from_numpy(w).cuda()
        out.fill_(float("nan"))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        sb.estimate_noalloc(d_xv.data_ptr(), d_w.data_ptr(), P, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        if call == 1:  # (the first call of a process may load the kernel's module; nothing after that may take memory)
            assert torch.cuda.mem_get_info()[0] >= free_before
        gx, gy, gt, gw = out.cpu().numpy()
        assert abs(gx - ex) <= TOL * max(1.0, abs(ex)), (call, gx, ex)
        assert abs(gy - ey) <= TOL * max(1.0, abs(ey)), (call, gy, ey)
        assert gt == et == float(xv[2, P // 3]), (call, gt, et)
        assert gw == 2.0
        # and against numpy directly (double accumulation, first maximum)
        assert abs(gx - xv[0].astype(np.float64).mean()) <= 1e-9 and abs(gy - xv[1].astype(np.float64).mean()) <= 1e-9
    f.close()


def test_noalloc_rejects_what_it_cannot_serve(sb):
    import torch
    torch.cuda.set_device(0)
    big = sb.NOALLOC_MAX_PARTICLES + 1
    buf = torch.zeros(16, dtype=torch.float32, device="cuda")
    out = torch.zeros(4, dtype=torch.float64, device="cuda")
    with pytest.raises(sb.SlamGpuError) as e:  # the size check comes before anything touches the (too small) buffers
        sb.estimate_noalloc(buf.data_ptr(), buf.data_ptr(), big, out.data_ptr())
    assert e.value.code == -1 and "SLAMGPU_NOALLOC_MAX_PARTICLES" in str(e.value)
    with pytest.raises(sb.SlamGpuError) as e:
        sb.estimate_noalloc(buf.data_ptr(), buf.data_ptr(), 0, out.data_ptr())
    assert e.value.code == -1
    host = np.zeros(16, np.float32)
    with pytest.raises(sb.SlamGpuError) as e:  # a host pointer is refused, not handed to the kernel
        sb.estimate_noalloc(host.ctypes.data, buf.data_ptr(), 4, out.data_ptr())
    assert e.value.code == -1 and "not device memory" in str(e.value)
    # the maximum itself is served
    P = sb.NOALLOC_MAX_PARTICLES
    xv = torch.ones((3, P), dtype=torch.float32, device="cuda")
    w = torch.ones(P, dtype=torch.float32, device="cuda")
    sb.estimate_noalloc(xv.data_ptr(), w.data_ptr(), P, out.data_ptr())
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [1.0, 1.0, 1.0, 1.0]
