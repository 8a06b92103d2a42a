"""viso_ba_eval_cost_static (include/viso_ba.h): the cost pass with its scratch in a file-scope static __device__ array,
for callers that cannot call cudaMalloc.  No counterpart in the reference,
so the checker is the existing entry point viso_ba_eval_cost on the same problem; the two sum in a
different order and agree to rounding (asserted at 1e-12 relative, far inside the 1e-5 cost tolerance of BASELINE).

Collected last on purpose (the file name).  The test passes alone in a fresh process
and after the other tests: it takes nothing from them but the session fixture that builds / loads the library."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12  # against viso_ba_eval_cost; the tests' cost tolerance against the oracle is 1e-5 (BASELINE north_star)
MAX_POSES = 4096  # VISO_BA_STATIC_MAX_POSES
OK, ERR_CAPACITY, ERR_ARG = 0, 1, 6  # viso_ba_status, include/viso_ba.h


def _device_arrays(torch, p):
    dev = torch.device("cuda:0")
    return {
        "poses": torch.from_numpy(np.ascontiguousarray(p["poses_init"], np.float64)).to(dev),
        "points": torch.from_numpy(np.ascontiguousarray(p["points_init"], np.float64)).to(dev),
        "edge_pose": torch.from_numpy(np.ascontiguousarray(p["edge_pose"], np.int32)).to(dev),
        "edge_point": torch.from_numpy(np.ascontiguousarray(p["edge_point"], np.int32)).to(dev),
        "edge_uv": torch.from_numpy(np.ascontiguousarray(p["edge_uv"], np.float32)).to(dev),
    }


def _static_cost(torch, lib, d, K, huber, out, n_poses=None):
    """one call through the C ABI on the default stream; returns (status, message)"""
    err = C.create_string_buffer(256)
    st = lib.viso_ba_eval_cost_static(
        None, int(d["poses"].shape[0] if n_poses is None else n_poses), d["poses"].data_ptr(), int(d["points"].shape[0]),
        d["points"].data_ptr(), int(d["edge_pose"].shape[0]), d["edge_pose"].data_ptr(), d["edge_point"].data_ptr(),
        d["edge_uv"].data_ptr(), K[0], K[1], K[2], K[3], float(huber), out.data_ptr(), err, len(err))
    return st, err.value.decode()


def _handle_cost(vb, p, huber, **cfg):
    ba = vb.BundleAdjuster(len(p["points_init"]), len(p["poses_init"]), len(p["edge_pose"]), huber_delta=huber, **cfg)
    ba.set_problem(p["poses_init"], p["points_init"], p["edge_pose"], p["edge_point"], p["edge_uv"])
    K = (ba.cfg.fx, ba.cfg.fy, ba.cfg.cx, ba.cfg.cy)
    cost = ba.eval_cost()
    ba.close()
    return cost, K


def test_static_scratch_cost_matches_the_handle_on_two_calls_in_a_row(vb):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    lib = vb.load_library()
    # call 1: 600 poses fill a good part of the scratch table; call 2: a smaller, different problem with another robust
    # kernel (whatever call 1 left in the array must not reach it); call 3: problem 1 again, same bits as call 1
    big = vb.synth_config(3, n_poses=600, n_points=20000, target_edges=160000, seed=0x5C7A7C01)
    small = vb.synth_config(2, seed=0x5C7A7C02)
    want_big, K_big = _handle_cost(vb, big, 2.0, path=2)
    want_small, K_small = _handle_cost(vb, small, 0.0)
    assert K_big == tuple(big["K"]) and K_small == tuple(small["K"])
    d_big, d_small = _device_arrays(torch, big), _device_arrays(torch, small)
    out = torch.full((3,), -1.0, dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()

    st, msg = _static_cost(torch, lib, d_big, K_big, 2.0, out[0:1])
    assert st == OK, msg
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]  # after the first call: its one-time module load is behind us
    st, msg = _static_cost(torch, lib, d_small, K_small, 0.0, out[1:2])
    assert st == OK, msg
    st, msg = _static_cost(torch, lib, d_big, K_big, 2.0, out[2:3])
    assert st == OK, msg
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before  # the driver handed out no memory: nothing was allocated
    got = out.cpu().numpy()
    assert np.isfinite(got).all() and want_big > 0 and want_small > 0
    assert got[0] == pytest.approx(want_big, rel=RTOL)
    assert got[1] == pytest.approx(want_small, rel=RTOL)
    assert got[2] == got[0]  # fixed summation shape: the same bits, whatever ran in between


def test_static_scratch_limits_and_bad_input(vb):
    import torch
    torch.cuda.set_device(0)
    lib = vb.load_library()
    p = vb.synth_config(1)
    d = _device_arrays(torch, p)
    out = torch.full((1,), -7.0, dtype=torch.float64, device="cuda:0")
    # one pose more than the scratch array holds: refused with a clear message, nothing enqueued, output untouched
    st, msg = _static_cost(torch, lib, d, p["K"], 0.0, out, n_poses=MAX_POSES + 1)
    assert st == ERR_CAPACITY
    assert str(MAX_POSES) in msg and "VISO_BA_STATIC_MAX_POSES" in msg
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == -7.0
    st, msg = _static_cost(torch, lib, d, p["K"], 0.0, out, n_poses=-1)
    assert st == ERR_ARG and msg
    # an edge that names a pose out of range is caught on the device: the cost is NaN, nothing is read out of bounds
    bad = dict(d)
    bad["edge_pose"] = d["edge_pose"].clone()
    bad["edge_pose"][3] = len(p["poses_init"])
    st, msg = _static_cost(torch, lib, bad, p["K"], 0.0, out)
    assert st == OK, msg
    torch.cuda.synchronize()
    assert np.isnan(float(out.cpu()[0]))
    # no edges: the cost is 0 and the stale partial sums of the call before are not read
    empty = dict(d)
    for k in ("edge_pose", "edge_point", "edge_uv"):
        empty[k] = d[k][:0]
    st, msg = _static_cost(torch, lib, empty, p["K"], 0.0, out)
    assert st == OK, msg
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == 0.0
    # and the array still serves a good call afterwards
    want, K = _handle_cost(vb, p, 0.0)
    st, msg = _static_cost(torch, lib, d, K, 0.0, out)
    assert st == OK, msg
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == pytest.approx(want, rel=RTOL)
