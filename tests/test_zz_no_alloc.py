"""umrs_scan_batch_reference_render_noalloc (include/umrs.h): the reference render with its
per-fan scratch in a device array compiled into the library -- no cudaMalloc on the call.
No counterpart in the reference, so the checker
is the existing entry point umrs_scan_batch_reference_render_device: same kernels, same bits.

Collected last on purpose; passes alone in a fresh process
and after the other tests.  Everything goes through the C-ABI; torch only holds the caller's
device buffers."""
import numpy as np
import pytest

from ut_multirobot_sim_b200 import synth

pytestmark = pytest.mark.gpu
D135 = np.float32(np.pi * 135.0 / 180.0)
N_RAYS = 1081


def _setup():
    import torch
    import ut_multirobot_sim_b200 as umrs
    from ut_multirobot_sim_b200 import _lib
    lib = _lib.load()
    segs = umrs.map_cleanup(synth.floorplan_segments(300, extent=60.0, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    poses = synth.sample_free_poses_fast(segs, 128, 60.0, seed=77)
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - D135, poses[:, 2] + D135], 1).astype(np.float32)
    # object lines: a few small closed squares near the first poses (occluders the trim pass must handle)
    sq = []
    for x, y in poses[:6, :2]:
        x, y = float(x) + 1.5, float(y) + 0.7
        sq += [[x, y, x + .4, y], [x + .4, y, x + .4, y + .4], [x + .4, y + .4, x, y + .4], [x, y + .4, x, y]]
    return torch, _lib, lib, vmap, fans, np.asarray(sq, np.float32)


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _both(torch, _lib, lib, vmap, fans, obj, stream=None):
    """(existing entry point, no-alloc entry point) on the same device inputs; outputs pre-filled with a canary."""
    n = len(fans)
    d_fans, d_obj = _dev(torch, fans), (_dev(torch, obj) if len(obj) else None)
    want = torch.full((n, N_RAYS), -7.0, dtype=torch.float32, device="cuda")
    got = torch.full((n, N_RAYS), -9.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    args = (vmap._h, d_obj.data_ptr() if d_obj is not None else None, len(obj), d_fans.data_ptr(), n, 100.0, N_RAYS)
    _lib.check(lib.umrs_scan_batch_reference_render_device(*args, want.data_ptr(), stream), "render_device")
    _lib.check(lib.umrs_scan_batch_reference_render_noalloc(*args, got.data_ptr(), stream), "render_noalloc")
    torch.cuda.synchronize()
    return want.cpu().numpy(), got.cpu().numpy()


def test_no_alloc_render_equals_the_existing_entry_point_on_two_calls_in_a_row():
    torch, _lib, lib, vmap, fans, obj = _setup()
    # set-up call: looks the scratch array up on this device, launches nothing
    _lib.check(lib.umrs_scan_batch_reference_render_noalloc(vmap._h, None, 0, None, 0, 100.0, N_RAYS, None, None), "setup")
    # call 1: the maximum batch, with object lines
    n_max = 64
    want1, got1 = _both(torch, _lib, lib, vmap, fans[:n_max], obj)
    assert np.isfinite(want1).all() and (want1 > 0).all()            # the canary is gone everywhere
    assert np.array_equal(want1.view(np.uint32), got1.view(np.uint32))
    # call 2, straight after: other fans, fewer of them, no object lines -- the shared array
    # still holds call 1's candidate lists, scenes and intervals
    want2, got2 = _both(torch, _lib, lib, vmap, fans[n_max:n_max + 37], obj[:0])
    assert np.array_equal(want2.view(np.uint32), got2.view(np.uint32))
    assert not np.array_equal(want2[:37], want1[:37])
    # and call 1's inputs once more give call 1's bits (nothing carried over from call 2)
    _, got3 = _both(torch, _lib, lib, vmap, fans[:n_max], obj)
    assert np.array_equal(got3.view(np.uint32), got1.view(np.uint32))


def test_no_alloc_render_rejects_batches_beyond_its_maximum():
    torch, _lib, lib, vmap, fans, obj = _setup()
    d_fans = _dev(torch, fans[:65])
    out = torch.full((65, N_RAYS), -9.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    st = lib.umrs_scan_batch_reference_render_noalloc(vmap._h, None, 0, d_fans.data_ptr(), 65, 100.0, N_RAYS,
                                                      out.data_ptr(), None)
    assert st == 1                                                    # UMRS_ERR_INVALID
    msg = lib.umrs_last_cuda_error().decode()
    assert "UMRS_NOALLOC_MAX_FANS" in msg and "65" in msg and "64" in msg
    torch.cuda.synchronize()
    assert (out == -9.0).all().item()                                 # nothing was launched
    # bad arguments
    assert lib.umrs_scan_batch_reference_render_noalloc(None, None, 0, d_fans.data_ptr(), 1, 100.0, N_RAYS, out.data_ptr(), None) == 1
    assert lib.umrs_scan_batch_reference_render_noalloc(vmap._h, None, 0, d_fans.data_ptr(), 1, 100.0, 0, out.data_ptr(), None) == 1
    assert lib.umrs_scan_batch_reference_render_noalloc(vmap._h, None, 0, None, 1, 100.0, N_RAYS, out.data_ptr(), None) == 1
    assert lib.umrs_scan_batch_reference_render_noalloc(vmap._h, None, 3, d_fans.data_ptr(), 1, 100.0, N_RAYS, out.data_ptr(), None) == 1


def test_no_alloc_render_runs_under_stream_capture():
    """Global-mode stream capture refuses cudaMalloc / cudaFree: the call is captured and the
    replayed graph gives the existing entry point's bits."""
    torch, _lib, lib, vmap, fans, obj = _setup()
    _lib.check(lib.umrs_scan_batch_reference_render_noalloc(vmap._h, None, 0, None, 0, 100.0, N_RAYS, None, None), "setup")
    n = 48
    want, _ = _both(torch, _lib, lib, vmap, fans[:n], obj)              # also loads the kernels before capture
    d_fans, d_obj = _dev(torch, fans[:n]), _dev(torch, obj)
    got = torch.full((n, N_RAYS), -9.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.graph(g, stream=s, capture_error_mode="global"):
        st = lib.umrs_scan_batch_reference_render_noalloc(vmap._h, d_obj.data_ptr(), len(obj), d_fans.data_ptr(), n, 100.0,
                                                          N_RAYS, got.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert st == 0, lib.umrs_last_cuda_error().decode()
    torch.cuda.synchronize()
    assert (got == -9.0).all().item()                                 # captured, not run
    for _ in range(2):
        got.fill_(-9.0)
        g.replay()
        torch.cuda.synchronize()
        assert np.array_equal(got.cpu().numpy().view(np.uint32), want.view(np.uint32))
