"""od_tdlt_fwd_batch_static -- the forward batch call that allocates nothing (plan records in
a static __device__ array, no context) -- against od_tdlt_fwd_batch, bit for bit, on two
calls in a row that share the array.  Collected last on purpose.
No counterpart in the reference."""
import numpy as np
import pytest

import daala_b200
from checker import make_maps

pytestmark = pytest.mark.gpu

KEY = daala_b200.OD_TDLT_KEYFRAME


def _run_both(torch, dev, pic_w, pic_h, decs, nframes, kind, flags, seed):
    rng = np.random.default_rng(seed)
    with daala_b200.TdltContext(pic_w, pic_h, decs=decs, device=0) as ctx:
        bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
        c = [torch.from_numpy(rng.integers(-2048, 2048, size=(nframes,) + s).astype(np.int32)).to(dev)
             for s in ctx.plane_shapes]
        maps = torch.from_numpy(bsize).to(dev)
        d_ctx = [torch.zeros_like(t) for t in c]
        d_sta = [torch.full_like(t, 12345) for t in c]
        st = torch.cuda.current_stream().cuda_stream
        cp = [t.data_ptr() for t in c]
        ctx.fwd_batch([t.data_ptr() for t in d_ctx], cp, maps.data_ptr(), nframes, flags, st)
        torch.cuda.synchronize()
        # everything the static call needs exists now: nothing may be allocated inside it
        before = torch.cuda.memory_allocated(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        rc = daala_b200.fwd_batch_static(pic_w, pic_h, decs, [t.data_ptr() for t in d_sta], cp,
                                         maps.data_ptr(), nframes, flags, st)
        torch.cuda.synchronize()
        assert rc == 0
        assert torch.cuda.memory_allocated(dev) == before
        for p in range(len(decs)):
            assert torch.equal(d_ctx[p], d_sta[p]), "plane %d differs (%s)" % (p, kind)
        return free_before, torch.cuda.mem_get_info(dev)[0]


def test_static_forward_matches_context_forward_twice_in_a_row():
    import torch
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    # first call: sets the array's address up; second call: other size, other maps, other
    # mode -- it finds the first call's records in the array and must not be misled by them
    _run_both(torch, dev, 352, 288, (0, 1, 1), 3, "quad", KEY, 11)
    f0, f1 = _run_both(torch, dev, 200, 136, (0, 1, 1), 2, "deep", 0, 12)
    assert f1 >= f0 - (2 << 20), "device memory shrank during the static call"
    # and once more at the first geometry, with different maps
    _run_both(torch, dev, 352, 288, (0, 1, 1), 3, "uniform1", KEY, 13)
    # lossless mode shares the array too
    _run_both(torch, dev, 352, 288, (0, 1, 1), 1, "uniform3", daala_b200.OD_TDLT_HAAR, 14)


def test_static_forward_rejects_batches_beyond_its_array(capfd):
    import torch
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    # 3840x2160 4:2:0 = 1530 regions per frame: 3 frames exceed the 4096 records
    t = torch.zeros(64, dtype=torch.int32, device=dev)
    ptrs = [t.data_ptr()] * 3
    rc = daala_b200.fwd_batch_static(3840, 2160, (0, 1, 1), ptrs, ptrs, t.data_ptr(), 3, KEY, 0)
    assert rc == -10    # OD_TDLT_EINVAL
    assert "OD_TDLT_STATIC_MAX_REGIONS" in capfd.readouterr().err
    torch.cuda.synchronize()
