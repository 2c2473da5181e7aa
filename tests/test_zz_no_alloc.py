"""obgpu_scan_offsets_static (include/obgpu.h): the offset scans of the symbolic phase with their scratch in a static device
array, against the existing entry points (obgpu_batch_upload_device + obgpu_symbolic + obgpu_get_offsets), through the C ABI.
Integer work: the comparison is exact.  Collected last on purpose."""
import numpy as np
import pytest

from openbeam_b200 import obgpu, workloads

pytestmark = pytest.mark.gpu


def _existing(first, count):
    ctx = obgpu.Context(0)
    ctx.upload(workloads.c3_batch(first, count), device=True)
    ctx.symbolic()
    offs = [ctx.offsets(w).copy() for w in (obgpu.OFF_DOF, obgpu.OFF_FREE, obgpu.OFF_BOUNDED)]
    ctx.close()
    return offs


def _static(offs):
    import torch
    S = len(offs[0]) - 1
    counts = np.zeros((S, 4), np.int32)
    for k in range(3):
        counts[:, k] = np.diff(offs[k])
    d_counts = torch.from_numpy(counts).cuda()
    outs = [torch.full((S + 1,), -7, dtype=torch.int64, device="cuda") for _ in range(3)]
    torch.cuda.synchronize()
    obgpu.scan_offsets_static(0, S, d_counts.data_ptr(), outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in outs]


def test_static_scratch_scan_matches_symbolic_twice_in_a_row():
    # two calls in a row share the scratch array: a large batch first (many chunks), then a small one whose partial sums
    # overwrite only the head of what the first call left behind
    for first, count in ((0, 5000), (7000, 1500)):
        ref = _existing(first, count)
        got = _static(ref)
        for r, g in zip(ref, got):
            assert r.dtype == g.dtype and np.array_equal(r, g)


def test_static_scratch_scan_rejects_sizes_beyond_the_maximum():
    import torch
    buf = torch.zeros(16, dtype=torch.int64, device="cuda")
    with pytest.raises(obgpu.ObgpuError) as ei:
        obgpu.scan_offsets_static(0, 4194304 + 1, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), buf.data_ptr())
    assert ei.value.status == obgpu.EINVAL and "maximum 4194304" in str(ei.value)
    with pytest.raises(obgpu.ObgpuError):
        obgpu.scan_offsets_static(0, 0, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), buf.data_ptr())
