"""doapp_distance_field_noalloc (static device scratch, no cudaMalloc) against the existing
doapp_grid_distance_field, through the C ABI, on two calls in a row.  Collected last on purpose.
Bar: bit-exact, as for every other kernel here (squared
distances are integers; the final sqrt * resolution is one correctly rounded operation each)."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def _reference_field(capi, occ, res):
    """The existing entry point: grid upload + doapp_grid_distance_field + download."""
    with capi.Context(k=1, m=1, n=1) as ctx:
        ctx.set_grid(occ, res)
        ctx.grid_distance_field()
        return ctx.grid_download(occ.shape, np.float32)


def _noalloc_field(capi, torch, occ, res):
    nz, ny, nx = occ.shape
    occ_d = torch.from_numpy(occ).cuda()
    out_d = torch.full((nz, ny, nx), -1.0, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream()
    torch.cuda.synchronize()
    capi.Context.distance_field_noalloc(occ_d.data_ptr(), nx, ny, nz, res, out_d.data_ptr(),
                                        stream.cuda_stream)
    stream.synchronize()
    return out_d.cpu().numpy()


def test_noalloc_distance_field_matches_existing_entry_point_twice(capi):
    import torch
    torch.cuda.set_device(0)
    # call 1: the largest grid the static scratch takes (128^3), a box + sparse random voxels;
    # call 2: a smaller, ragged grid with other contents -- the scratch still holds call 1's
    # planes, which must not leak into the result
    occ1 = synth.box_grid(128, 0.01, fill=0.0005)
    rng = np.random.default_rng(5)
    occ2 = (rng.random((37, 50, 61)) < 0.002).astype(np.uint8)
    for occ, res in ((occ1, 0.01), (occ2, 0.02)):
        want = _reference_field(capi, occ, res)
        got = _noalloc_field(capi, torch, occ, res)
        assert np.isfinite(want).all() and (want > 0).any()
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    # an empty grid: +inf everywhere, on both paths
    empty = np.zeros((8, 8, 8), np.uint8)
    assert np.array_equal(_noalloc_field(capi, torch, empty, 0.02), _reference_field(capi, empty, 0.02))


def test_noalloc_rejects_sizes_beyond_the_static_scratch(capi):
    import torch
    torch.cuda.set_device(0)
    buf = torch.zeros(16, dtype=torch.uint8, device="cuda")
    out = torch.zeros(16, dtype=torch.float32, device="cuda")
    with pytest.raises(capi.DoappError) as e:  # 129 * 128 * 128 cells > DOAPP_NOALLOC_MAX_CELLS
        capi.Context.distance_field_noalloc(buf.data_ptr(), 129, 128, 128, 0.01, out.data_ptr())
    assert "exceed the static scratch" in str(e.value)
    with pytest.raises(capi.DoappError):
        capi.Context.distance_field_noalloc(0, 8, 8, 8, 0.01, out.data_ptr())
