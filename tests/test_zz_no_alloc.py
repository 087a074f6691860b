"""xrt_morton_codes_noalloc (include/xrt_b200.h): the allocation-free variant of the tree build's first stage, through
the C ABI, against the existing allocating entry point xrt_bvh_build_standalone.  Bit-exact (integer codes; the scene
box is a min / max).  Collected last on purpose (file name)."""
import numpy as np
import pytest

import xrtailor_b200 as X

MAX_FACES = 262144  # XRT_NOALLOC_MAX_FACES


def grid_mesh(n, side, origin, seed):
    """n x n vertices, 2 (n-1)^2 triangles, jittered so Morton codes are not degenerate"""
    rng = np.random.default_rng(seed)
    u = np.linspace(0.0, side, n, dtype=np.float32)
    gx, gz = np.meshgrid(u, u, indexing="ij")
    pos = np.stack([gx, 0.05 * np.sin(7.0 * gx) * np.cos(5.0 * gz), gz], axis=-1).reshape(-1, 3).astype(np.float32)
    pos += rng.uniform(-1e-3, 1e-3, pos.shape).astype(np.float32) + np.asarray(origin, np.float32)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    a = (i * n + j).ravel()
    tri = np.concatenate([np.stack([a, a + 1, a + n], 1), np.stack([a + 1, a + n + 1, a + n], 1)]).astype(np.uint32)
    return pos, tri


def noalloc_codes(torch, pos, tri, fast_div, stream=0):
    dev = torch.device("cuda:0")
    xyzw = np.zeros((len(pos), 4), np.float32)
    xyzw[:, :3] = pos
    d_x = torch.from_numpy(xyzw).to(dev)
    d_t = torch.from_numpy(tri.astype(np.int64)).to(dev).to(torch.int32).contiguous()  # same 32-bit pattern as uint32
    d_m = torch.full((len(tri),), -1, dtype=torch.int32, device=dev)
    d_b = torch.zeros(6, dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    X.morton_codes_noalloc(d_x.data_ptr(), len(pos), d_t.data_ptr(), len(tri), d_m.data_ptr(), d_b.data_ptr(),
                           fast_div=fast_div, stream=stream)
    torch.cuda.synchronize()
    return d_m.cpu().numpy().view(np.uint32), d_b.cpu().numpy()


def check_against_existing(torch, pos, tri, fast_div):
    codes, box = noalloc_codes(torch, pos, tri, fast_div)
    ref = X.bvh_build_standalone(pos, tri, fast_div=fast_div)  # allocates, builds the whole tree, sorts the codes
    assert np.array_equal(ref["morton"], codes[ref["object"]])
    assert np.array_equal(np.sort(codes), ref["morton"])
    used = pos[np.unique(tri)]
    assert np.array_equal(box[:3], used.min(0)) and np.array_equal(box[3:], used.max(0))
    return codes


@pytest.mark.gpu
def test_noalloc_morton_codes_equal_the_allocating_entry_two_calls_in_a_row():
    import torch
    torch.cuda.set_device(0)
    big = grid_mesh(120, 1.0, (0.0, 1.0, 0.0), 20261104)     # 28 322 faces
    small = grid_mesh(37, 0.25, (-2.0, 0.3, 4.0), 7)         # 2 592 faces, a different scene box
    for fast_div in (0, 1):
        a = check_against_existing(torch, *big, fast_div)
        b = check_against_existing(torch, *small, fast_div)  # the shared scratch still holds the big mesh's boxes
        a2 = check_against_existing(torch, *big, fast_div)
        assert np.array_equal(a, a2) and len(b) == len(small[1])


@pytest.mark.gpu
def test_noalloc_on_a_side_stream_and_device_memory_unchanged():
    import torch
    torch.cuda.set_device(0)
    pos, tri = grid_mesh(64, 0.8, (0.5, 0.0, -0.5), 3)
    st = torch.cuda.Stream()
    first, _ = noalloc_codes(torch, pos, tri, 1, stream=st.cuda_stream)  # also the per-device address lookup
    dev = torch.device("cuda:0")
    xyzw = np.zeros((len(pos), 4), np.float32)
    xyzw[:, :3] = pos
    d_x = torch.from_numpy(xyzw).to(dev)
    d_t = torch.from_numpy(tri.astype(np.int64)).to(dev).to(torch.int32).contiguous()
    d_m = torch.zeros(len(tri), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(2):
        X.morton_codes_noalloc(d_x.data_ptr(), len(pos), d_t.data_ptr(), len(tri), d_m.data_ptr(), 0, fast_div=1,
                               stream=st.cuda_stream)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "the call must not allocate device memory"
    assert np.array_equal(d_m.cpu().numpy().view(np.uint32), first)


@pytest.mark.gpu
def test_noalloc_rejects_sizes_beyond_the_static_scratch_and_host_pointers():
    import torch
    torch.cuda.set_device(0)
    d = torch.zeros(64, dtype=torch.float32, device="cuda:0")
    with pytest.raises(X.XrtError) as e:
        X.morton_codes_noalloc(d.data_ptr(), 16, d.data_ptr(), MAX_FACES + 1, d.data_ptr())
    assert e.value.code == 8100 and "XRT_NOALLOC_MAX_FACES" in str(e.value) and str(MAX_FACES) in str(e.value)
    host = np.zeros(64, np.float32)
    with pytest.raises(X.XrtError) as e:
        X.morton_codes_noalloc(host.ctypes.data, 16, d.data_ptr(), 4, d.data_ptr())
    assert e.value.code == 8100 and "device memory" in str(e.value)
