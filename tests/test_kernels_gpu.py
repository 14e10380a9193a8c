"""GPU tests of single C-ABI entry points (sort, ranges, key emission) and of the operator's
edge cases (empty frustum, empty key list, rgb_only, hook-less backward, full-size run)."""
import ctypes as C

import numpy as np
import pytest
import torch

from taichi_3d_gaussian_splatting_b200.synthetic import make_scene

import util

pytestmark = pytest.mark.gpu


def _lib():
    from taichi_3d_gaussian_splatting_b200 import _capi
    return _capi


def _sort_pairs(keys_np, depth_bits, tile_bits):
    capi = _lib()
    k = keys_np.shape[0]
    dev = torch.device("cuda")
    keys = torch.from_numpy(keys_np.copy()).to(dev)
    vals = torch.arange(k, dtype=torch.int32, device=dev)
    keys_alt, vals_alt = torch.empty_like(keys), torch.empty_like(vals)
    ws = torch.empty(int(capi.lib.gs_sort_workspace_bytes(k)), dtype=torch.uint8, device=dev)
    in_alt = C.c_int(0)
    st = torch.cuda.current_stream().cuda_stream
    capi.check(capi.lib.gs_sort_pairs(keys.data_ptr(), vals.data_ptr(), keys_alt.data_ptr(),
                                      vals_alt.data_ptr(), k, depth_bits, tile_bits, ws.data_ptr(),
                                      C.byref(in_alt), st), "gs_sort_pairs")
    torch.cuda.synchronize()
    rk, rv = (keys_alt, vals_alt) if in_alt.value else (keys, vals)
    return rk.cpu().numpy(), rv.cpu().numpy()


@pytest.mark.parametrize("k", [1, 2, 31, 4095, 4096, 4097, 12289, 300001, 3000000])
@pytest.mark.parametrize("bits", [(17, 13), (15, 12), (8, 1), (5, 3)])
def test_sort_significant_bits_stable(k, bits):
    depth_bits, tile_bits = bits
    rng = np.random.default_rng(k * 31 + depth_bits)
    # few distinct depths -> many ties, so stability is really exercised
    depth = rng.integers(0, min(1 << depth_bits, 50), k, dtype=np.int64)
    tile = rng.integers(0, 1 << tile_bits, k, dtype=np.int64)
    keys = depth + (tile << 32)
    sk, sv = _sort_pairs(keys, depth_bits, tile_bits)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(sk, keys[order])
    assert np.array_equal(sv, order.astype(np.int32))


@pytest.mark.parametrize("k", [5, 4097, 100000])
def test_sort_full_signed_64bit(k):
    rng = np.random.default_rng(k)
    keys = rng.integers(-2**40, 2**40, k, dtype=np.int64)
    keys[::7] = keys[0]
    sk, sv = _sort_pairs(keys, 32, 32)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(sk, keys[order])
    assert np.array_equal(sv, order.astype(np.int32))


@pytest.mark.parametrize("k", [1, 2, 33, 4096, 4097, 250001, 3000000])
@pytest.mark.parametrize("bits", [13, 8, 16, 3, 17])
def test_sort_pairs32_stable(k, bits):
    """(u32 tile id, i32 payload) pairs by the low `bits` bits, ties in input order."""
    capi = _lib()
    rng = np.random.default_rng(k * 7 + bits)
    keys_np = rng.integers(0, min(1 << bits, 9000), k, dtype=np.int64).astype(np.uint32)
    dev = torch.device("cuda")
    keys = torch.from_numpy(keys_np.view(np.int32).copy()).to(dev)
    vals = torch.arange(k, dtype=torch.int32, device=dev)
    keys_alt, vals_alt = torch.empty_like(keys), torch.empty_like(vals)
    ws = torch.empty(int(capi.lib.gs_sort_workspace_bytes(k)), dtype=torch.uint8, device=dev)
    in_alt = C.c_int(0)
    capi.check(capi.lib.gs_sort_pairs32(keys.data_ptr(), vals.data_ptr(), keys_alt.data_ptr(),
                                        vals_alt.data_ptr(), k, bits, ws.data_ptr(), None, C.byref(in_alt),
                                        torch.cuda.current_stream().cuda_stream), "gs_sort_pairs32")
    torch.cuda.synchronize()
    rk, rv = (keys_alt, vals_alt) if in_alt.value else (keys, vals)
    order = np.argsort(keys_np, kind="stable")
    assert np.array_equal(rk.cpu().numpy().view(np.uint32), keys_np[order])
    assert np.array_equal(rv.cpu().numpy(), order.astype(np.int32))


@pytest.mark.parametrize("m", [1, 31, 4097, 70001, 1000000])
@pytest.mark.parametrize("bits", [17, 8, 24, 31, 1])
def test_sort_points_by_depth_is_a_stable_argsort(m, bits):
    capi = _lib()
    rng = np.random.default_rng(m + bits)
    depth = rng.integers(0, min(1 << bits, 5000) if bits < 24 else (1 << bits), m, dtype=np.int64)
    depth = depth.astype(np.int32)
    dev = torch.device("cuda")
    d = torch.from_numpy(depth).to(dev)
    order = torch.empty(m, dtype=torch.int32, device=dev)
    ws = torch.empty(int(capi.lib.gs_point_sort_workspace_bytes(m)), dtype=torch.uint8, device=dev)
    capi.check(capi.lib.gs_sort_points_by_depth(d.data_ptr(), m, None, bits, order.data_ptr(), ws.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream),
               "gs_sort_points_by_depth")
    torch.cuda.synchronize()
    assert np.array_equal(d.cpu().numpy(), depth)  # input untouched
    want = np.argsort(depth, kind="stable").astype(np.int32)
    assert np.array_equal(order.cpu().numpy(), want)
    # capacity form: the count is read on the device, buffers sized for a larger capacity
    cap = m + 5000
    d2 = torch.cat([d, torch.full((cap - m,), 7, dtype=torch.int32, device=dev)])
    order2 = torch.full((cap,), -1, dtype=torch.int32, device=dev)
    ws2 = torch.empty(int(capi.lib.gs_point_sort_workspace_bytes(cap)), dtype=torch.uint8, device=dev)
    count = torch.tensor([m, 0], dtype=torch.int64, device=dev)
    capi.check(capi.lib.gs_sort_points_by_depth(d2.data_ptr(), cap, count.data_ptr(), bits, order2.data_ptr(),
                                                ws2.data_ptr(), torch.cuda.current_stream().cuda_stream),
               "gs_sort_points_by_depth")
    torch.cuda.synchronize()
    assert np.array_equal(order2[:m].cpu().numpy(), want)


def test_emit_tile_keys_follows_the_given_order():
    """Tile ids / payload contiguous per point, points in `order`; whole-image, single-tile and
    zero-tile boxes mixed; tile ranges from the sorted ids."""
    capi = _lib()
    W, H = 320, 208
    tw, th = W // 16, H // 16
    rng = np.random.default_rng(9)
    m = 3000
    tu0 = rng.integers(0, tw + 1, m); tu1 = np.minimum(tu0 + rng.integers(0, 6, m), tw)
    tv0 = rng.integers(0, th + 1, m); tv1 = np.minimum(tv0 + rng.integers(0, 6, m), th)
    tu0[::13], tu1[::13], tv0[::13], tv1[::13] = 0, tw, 0, th
    bbox = np.stack([tu0, tu1, tv0, tv1], 1).astype(np.uint16)
    order_np = rng.permutation(m).astype(np.int32)
    ids_ref, pay_ref = [], []
    for r in order_np:
        for u in range(bbox[r, 0], bbox[r, 1]):
            for v in range(bbox[r, 2], bbox[r, 3]):
                ids_ref.append(u + v * tw)
                pay_ref.append(r)
    k = len(ids_ref)
    dev = torch.device("cuda")
    d_bbox = torch.from_numpy(bbox.view(np.int16)).to(dev)
    d_order = torch.from_numpy(order_np).to(dev)
    ids = torch.empty(k, dtype=torch.int32, device=dev)
    pay = torch.empty(k, dtype=torch.int32, device=dev)
    ws = torch.empty(int(capi.lib.gs_emit_workspace_bytes(m)), dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    tile_bits = max(int(tw * th - 1).bit_length(), 1)
    hist = torch.full((4 * 256,), -1, dtype=torch.int32, device=dev)   # the call zeroes it
    capi.check(capi.lib.gs_emit_tile_keys(d_order.data_ptr(), d_bbox.data_ptr(), m, k, W, ids.data_ptr(),
                                          pay.data_ptr(), ws.data_ptr(), hist.data_ptr(), tile_bits, st),
               "gs_emit_tile_keys")
    torch.cuda.synchronize()
    assert np.array_equal(ids.cpu().numpy(), np.asarray(ids_ref, np.int32))
    assert np.array_equal(pay.cpu().numpy(), np.asarray(pay_ref, np.int32))
    # the digit histogram handed to the pair sort: exact counts per 8-bit digit, unused digits zero
    h = hist.cpu().numpy().reshape(4, 256)
    ref_ids = np.asarray(ids_ref, np.uint32)
    for p in range(4):
        want = np.bincount((ref_ids >> (8 * p)) & 255, minlength=256) if 8 * p < tile_bits else np.zeros(256, int)
        assert np.array_equal(h[p], want), p
    # sorting with that histogram gives the same result as sorting without it
    outs = []
    for use_hist in (False, True):
        k2 = torch.cat([ids, torch.empty_like(ids)])
        v2 = torch.cat([pay, torch.empty_like(pay)])
        sws = torch.empty(int(capi.lib.gs_sort_workspace_bytes(k)), dtype=torch.uint8, device=dev)
        in_alt = C.c_int(0)
        capi.check(capi.lib.gs_sort_pairs32(k2.data_ptr(), v2.data_ptr(), k2.data_ptr() + 4 * k,
                                            v2.data_ptr() + 4 * k, k, tile_bits, sws.data_ptr(),
                                            hist.data_ptr() if use_hist else None, C.byref(in_alt), st),
                   "gs_sort_pairs32")
        torch.cuda.synchronize()
        half = slice(k, 2 * k) if in_alt.value else slice(0, k)
        outs.append((k2[half].cpu().numpy(), v2[half].cpu().numpy()))
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    order_ref = np.argsort(ref_ids, kind="stable")
    assert np.array_equal(outs[1][1], np.asarray(pay_ref, np.int32)[order_ref])
    # without a histogram buffer the emission is unchanged
    ids_b = torch.empty_like(ids)
    capi.check(capi.lib.gs_emit_tile_keys(d_order.data_ptr(), d_bbox.data_ptr(), m, k, W, ids_b.data_ptr(),
                                          pay.data_ptr(), ws.data_ptr(), None, 0, st), "gs_emit_tile_keys")
    torch.cuda.synchronize()
    assert torch.equal(ids_b, ids)
    sorted_ids = torch.sort(ids.to(torch.int64)).values.to(torch.int32)
    start = torch.empty(tw * th, dtype=torch.int32, device=dev)
    end = torch.empty(tw * th, dtype=torch.int32, device=dev)
    capi.check(capi.lib.gs_tile_ranges32(sorted_ids.data_ptr(), k, start.data_ptr(), end.data_ptr(),
                                         tw * th, st), "gs_tile_ranges32")
    torch.cuda.synchronize()
    counts = np.bincount(np.asarray(ids_ref), minlength=tw * th)
    s_np, e_np = start.cpu().numpy(), end.cpu().numpy()
    assert np.array_equal(e_np - s_np, counts)
    nz = counts > 0
    assert np.array_equal(s_np[nz], (np.cumsum(counts) - counts)[nz]) and not s_np[~nz].any()


def _bin_tiles_case(W, H, m, rng, wide_every=13, empty_every=7, max_span=6):
    tw, th = W // 16, H // 16
    tu0 = rng.integers(0, tw + 1, m); tu1 = np.minimum(tu0 + rng.integers(0, max_span, m), tw)
    tv0 = rng.integers(0, th + 1, m); tv1 = np.minimum(tv0 + rng.integers(0, max_span, m), th)
    if wide_every:
        tu0[::wide_every], tu1[::wide_every], tv0[::wide_every], tv1[::wide_every] = 0, tw, 0, th
    if empty_every:
        tu1[3::empty_every] = tu0[3::empty_every]
    return np.stack([tu0, tu1, tv0, tv1], 1).astype(np.uint16)


@pytest.mark.parametrize("shape", [(320, 208, 3000), (16, 16, 5000), (4096, 16, 2500), (16, 4096, 2500),
                                   (1920, 1088, 200000), (64, 48, 1), (976, 544, 1023), (976, 544, 1025)])
def test_bin_tiles_equals_stable_sort_by_tile_id(shape):
    """gs_bin_tiles (column split, then row split of the depth-ordered points) gives the payload order
    and the tile ranges of a stable sort of the emitted (tile id, payload) pairs by tile id -- the
    order R.py:947-950 produces with stable ties -- including whole-image, empty and clipped boxes."""
    capi = _lib()
    W, H, m = shape
    tw, th = W // 16, H // 16
    rng = np.random.default_rng(W * 31 + H + m)
    bbox = _bin_tiles_case(W, H, m, rng, wide_every=13 if m <= 5000 else 997,
                           max_span=6 if m <= 5000 else 4)
    order_np = rng.permutation(m).astype(np.int32)
    b = bbox[order_np].astype(np.int64)
    wdt, hgt = np.maximum(b[:, 1] - b[:, 0], 0), np.maximum(b[:, 3] - b[:, 2], 0)
    cnt = wdt * hgt
    k = int(cnt.sum())
    rep = np.repeat(np.arange(m), cnt)
    within = np.arange(k) - np.repeat(np.cumsum(cnt) - cnt, cnt)
    hh = np.maximum(hgt[rep], 1)
    ids_ref = (b[rep, 0] + within // hh) + (b[rep, 2] + within % hh) * tw   # emission order of R.py:161-172
    pay_ref = order_np[rep]
    srt = np.argsort(ids_ref, kind="stable")
    want_pay = pay_ref[srt]
    counts = np.bincount(ids_ref, minlength=tw * th)
    dev = torch.device("cuda")
    d_bbox = torch.from_numpy(bbox.view(np.int16)).to(dev)
    d_order = torch.from_numpy(order_np).to(dev)
    cap = int(capi.lib.gs_bin_entries_capacity(m, k, W))
    assert cap == min(k, m * tw)
    entries = torch.empty(max(cap, 1), dtype=torch.int64, device=dev)
    pay = torch.full((max(k, 1),), -7, dtype=torch.int32, device=dev)
    ws = torch.empty(int(capi.lib.gs_bin_workspace_bytes(m, k, W, H)), dtype=torch.uint8, device=dev)
    start = torch.full((tw * th,), -1, dtype=torch.int32, device=dev)
    end = torch.full((tw * th,), -1, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):   # the workspace is reusable without clearing by the caller
        capi.check(capi.lib.gs_bin_tiles(d_order.data_ptr(), d_bbox.data_ptr(), m, k, W, H, entries.data_ptr(),
                                         pay.data_ptr(), start.data_ptr(), end.data_ptr(), ws.data_ptr(), st),
                   "gs_bin_tiles")
        torch.cuda.synchronize()
        assert np.array_equal(pay[:k].cpu().numpy(), want_pay)
        s_np, e_np = start.cpu().numpy(), end.cpu().numpy()
        assert np.array_equal(e_np - s_np, counts)
        nz = counts > 0
        assert np.array_equal(s_np[nz], (np.cumsum(counts) - counts)[nz]) and not s_np[~nz].any()


def test_bin_tiles_empty_and_capacity():
    capi = _lib()
    dev = torch.device("cuda")
    start = torch.full((12,), -1, dtype=torch.int32, device=dev)
    end = torch.full((12,), -1, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    capi.check(capi.lib.gs_bin_tiles(None, None, 0, 0, 64, 48, None, None, start.data_ptr(), end.data_ptr(),
                                     None, st), "gs_bin_tiles")
    torch.cuda.synchronize()
    assert not start.any() and not end.any()
    assert capi.lib.gs_bin_tiles(None, None, 0, 0, 16 * 257, 16, None, None, start.data_ptr(),
                                 end.data_ptr(), None, st) == -2  # GS_E_CAPACITY
    assert capi.lib.gs_bin_workspace_bytes(10, 10, 16 * 257, 16) == -2  # GS_E_CAPACITY


def test_tile_ranges_reference_golden_vector():
    """tests/GaussianPointCloudRasterisation_test.py:18-51 of the reference, same call."""
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import find_tile_start_and_end
    keys = torch.tensor([0x100000000, 0x100000001, 0x200000000, 0x200000001, 0x200000002,
                         0x300000000, 0x300000001], dtype=torch.int64, device="cuda")
    start = torch.zeros(4, dtype=torch.int32, device="cuda")
    end = torch.zeros(4, dtype=torch.int32, device="cuda")
    find_tile_start_and_end(keys, start, end)
    assert start.tolist() == [0, 0, 2, 5]
    assert end.tolist() == [0, 2, 5, 7]


def test_emit_keys_large_and_empty_boxes(oracle32):
    """Splats covering the whole image, single-tile splats and zero-tile splats mixed."""
    capi = _lib()
    W, H = 320, 208
    tw, th = W // 16, H // 16
    rng = np.random.default_rng(5)
    m = 1500
    uv = np.stack([rng.uniform(-40, W + 40, m), rng.uniform(-40, H + 40, m)], 1).astype(np.float32)
    radii = np.exp(rng.uniform(np.log(0.5), np.log(400.0), m)).astype(np.float32)
    radii[::11] = 5000.0
    pic = np.zeros((m, 3), np.float32)
    pic[:, 2] = rng.uniform(1, 900, m)
    counts = oracle32.tile_counts(uv, radii, W, H)
    assert (counts == 0).any() and (counts == tw * th).any()
    offsets, total = oracle32.exclusive_scan(counts)
    keys_ref, pay_ref = oracle32.emit_keys(uv, pic, radii, offsets, total, W, H, 100.0)
    # bbox / depth key exactly as the projection kernel would hand them over
    bbox = np.zeros((m, 4), np.uint16)
    for i in range(m):
        r = max(radii[i], np.float32(1.0))
        mn_u, mx_u = max(np.float32(0), uv[i, 0] - r), uv[i, 0] + r
        mn_v, mx_v = max(np.float32(0), uv[i, 1] - r), uv[i, 1] + r
        tu0 = min(int(np.floor(mn_u / 16)), tw)
        tu1 = min(max(int(np.floor(mx_u / 16)) + 1, tu0 + 1), tw)
        tv0 = min(int(np.floor(mn_v / 16)), th)
        tv1 = min(max(int(np.floor(mx_v / 16)) + 1, tv0 + 1), th)
        bbox[i] = (tu0, tu1, tv0, tv1)
    assert np.array_equal((bbox[:, 1].astype(int) - bbox[:, 0]) * (bbox[:, 3].astype(int) - bbox[:, 2]),
                          counts)
    depth_key = (pic[:, 2] * np.float32(100.0)).astype(np.int32)
    dev = torch.device("cuda")
    d_bbox = torch.from_numpy(bbox.view(np.int16)).to(dev)
    d_depth = torch.from_numpy(depth_key).to(dev)
    d_off = torch.from_numpy(offsets).to(dev)
    keys = torch.empty(total, dtype=torch.int64, device=dev)
    pay = torch.empty(total, dtype=torch.int32, device=dev)
    capi.check(capi.lib.gs_emit_keys(d_bbox.data_ptr(), d_depth.data_ptr(), d_off.data_ptr(), m,
                                     total, W, keys.data_ptr(), pay.data_ptr(),
                                     torch.cuda.current_stream().cuda_stream), "gs_emit_keys")
    torch.cuda.synchronize()
    assert np.array_equal(keys.cpu().numpy(), keys_ref)
    assert np.array_equal(pay.cpu().numpy(), pay_ref)


def _module(near=0.8, far=1000.0, scale=100.0, rgb_only=False, hook=None):
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation
    return GaussianPointCloudRasterisation(
        GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig(
            near_plane=near, far_plane=far, depth_to_sort_key_scale=scale, rgb_only=rgb_only),
        backward_valid_point_hook=hook)


def _call(ras, scene, t, band=2):
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation
    cam = CameraInfo(t["camera_intrinsics"], scene.height, scene.width, 0)
    return ras(GaussianPointCloudRasterisation.GaussianPointCloudRasterisationInput(
        point_cloud=t["point_cloud"], point_cloud_features=t["point_cloud_features"],
        point_object_id=t["point_object_id"], point_invalid_mask=t["point_invalid_mask"],
        camera_info=cam, q_pointcloud_camera=t["q_pointcloud_camera"],
        t_pointcloud_camera=t["t_pointcloud_camera"], color_max_sh_band=band))


def test_empty_frustum_gives_zero_image_and_zero_grads():
    scene = make_scene(500, 64, 48, seed=9, scale_factor=0.05)
    scene.point_invalid_mask[:] = 1
    t = util.scene_tensors(scene, requires_grad=True)
    hooked = {}
    ras = _module(hook=lambda h: hooked.update(h=h))
    image, depth, count = _call(ras, scene, t)
    assert image.shape == (48, 64, 3) and depth.shape == (48, 64) and count.dtype == torch.int32
    assert float(image.abs().max()) == 0 and int(count.max()) == 0
    image.sum().backward()
    assert float(t["point_cloud"].grad.abs().max()) == 0
    assert float(t["point_cloud_features"].grad.abs().max()) == 0
    assert hooked["h"].point_id_in_camera_list.shape[0] == 0
    assert ras.last_forward_info["num_points_in_camera"] == 0


def test_visible_points_with_zero_tiles(oracle32):
    """Points inside the 48 px margin right of / below the image are visible but touch no tile
    (R.py:94-102): M > 0, K == 0."""
    scene = make_scene(200, 64, 64, seed=10, scale_factor=0.001)
    f = 0.6 * 64
    z = scene.point_cloud[:, 2]
    scene.point_cloud[:, 0] = (64 + 20 - 32) * z / f   # u = W + 20
    scene.point_invalid_mask[:] = 0
    ref = util.oracle_forward(oracle32, scene)
    assert len(ref["point_id_in_camera_list"]) == 200 and len(ref["point_in_camera_sort_key"]) == 0
    got = util.cuda_forward_stages(scene)
    assert not util.mismatch_report(ref, got)


def test_rgb_only_matches_image(oracle32):
    scene = make_scene(2000, 64, 64, seed=11, scale_factor=0.05)
    ref = util.oracle_forward(oracle32, scene, rgb_only=True)
    got = util.cuda_forward_stages(scene, rgb_only=True)
    assert np.array_equal(ref["rasterized_image"], got["rasterized_image"])
    assert not got["rasterized_depth"].any() and not got["pixel_valid_point_count"].any()


def test_reference_two_points_case(oracle32):
    """tests/GaussianPointCloudRasterisation_test.py:152-205 of the reference: two points on the
    optical axis, the nearer one invalid, 16x16 image, near 0, far 10."""
    from taichi_3d_gaussian_splatting_b200.synthetic import SyntheticScene
    feat = np.zeros((2, 56), np.float32)
    feat[:, 3] = 1.0
    feat[0, 4:7] = 1.0
    feat[1, 4:7] = 4.0
    feat[:, 8] = 5.0
    feat[:, 24] = 1.0
    feat[:, 40] = 1.0
    scene = SyntheticScene(np.array([[0, 0, 1], [0, 0, 2]], np.float32), feat,
                           np.array([1, 0], np.int8), np.zeros(2, np.int32),
                           np.array([[1, 0, 8], [0, 1, 8], [0, 0, 1]], np.float32),
                           np.array([[0, 0, 0, 1]], np.float32), np.zeros((1, 3), np.float32), 16, 16)
    ref = util.oracle_forward(oracle32, scene, 0.0, 10.0, 100.0)
    got = util.cuda_forward_stages(scene, 0.0, 10.0, 100.0)
    assert list(ref["point_id_in_camera_list"]) == [1]
    assert not util.mismatch_report(ref, got)
    assert got["rasterized_image"].max() > 0.1


def test_inference_under_no_grad_and_input_validation():
    scene = make_scene(3000, 64, 64, seed=12, scale_factor=0.05)
    t = util.scene_tensors(scene)
    ras = _module()
    with torch.no_grad():
        image, depth, count = _call(ras, scene, t)
    assert image.requires_grad is False and torch.isfinite(image).all()
    bad = dict(t)
    bad["point_cloud_features"] = t["point_cloud_features"].double()
    with pytest.raises(ValueError):
        _call(ras, scene, bad)
    scene_bad = make_scene(10, 64, 64, seed=1)
    scene_bad.width = 60
    with pytest.raises(AssertionError):  # R.py:1193-1194
        _call(ras, scene_bad, util.scene_tensors(scene_bad))
    cpu = util.scene_tensors(scene, device="cpu")
    with pytest.raises(Exception):
        _call(ras, scene, cpu)


def test_training_loop_reduces_loss():
    """Behavioural pin of reference test_backward_coverage (T_R.py:284-351), shortened: Adam on a
    32x32 synthetic target, loss must go down."""
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation
    torch.manual_seed(0)
    dev = "cuda"
    n = 10000
    target = torch.zeros(32, 32, 3, device=dev)
    target[:5, :2, 0] = 1.0
    target[:5, :2, 1] = 0.7
    target[8:24, 8:24, 0] = 0.5
    target[8:24, 8:24, 1] = 0.7
    target[20:28, 20:28, 0] = 0.8
    target[20:28, 20:28, 1] = 0.1
    xyz = torch.nn.Parameter((torch.rand(n, 3, device=dev) - 0.5) * 3)
    tmp = torch.rand(n, 56, device=dev)
    tmp[:, 4:7] = -4.60517018599
    tmp[:, 7] = 0.5
    feat = torch.nn.Parameter(tmp)
    cam = CameraInfo(torch.tensor([[32., 0, 16], [0, 32., 16], [0, 0, 1]], device=dev), 32, 32, 0)
    q = torch.tensor([[0., 0, 0, 1]], device=dev)
    tt = torch.tensor([[0., 0, 2.0]], device=dev)  # T_pointcloud_camera: camera 2 units behind
    ras = GaussianPointCloudRasterisation(
        GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig(near_plane=1., far_plane=10.))
    opt = torch.optim.Adam([xyz, feat], lr=0.001)
    inv = torch.zeros(n, dtype=torch.int8, device=dev)
    oid = torch.zeros(n, dtype=torch.int32, device=dev)
    losses = []
    for it in range(300):
        opt.zero_grad()
        image, _, _ = ras(GaussianPointCloudRasterisation.GaussianPointCloudRasterisationInput(
            point_cloud=xyz, point_cloud_features=feat, point_object_id=oid, point_invalid_mask=inv,
            camera_info=cam, q_pointcloud_camera=q, t_pointcloud_camera=-tt,
            color_max_sh_band=it // 100))
        loss = ((image - target) ** 2).sum()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert np.isfinite(losses).all()
    assert losses[-1] < 0.7 * losses[0], (losses[0], losses[-1])


FULL_SIZE_CASES = {
    # name: (bench workload, (near, far, key scale), tile-row band of the gradient check or None)
    "c2": ("c2_1m_1920x1088", (0.8, 1000.0, 100.0), None),
    "c3": ("c3_430k_976x544", (0.8, 1000.0, 100.0), None),
    # the Truck run's rasteriser values (reference config/tat_truck_every_8_test.yaml:44-47)
    "c3_truck_cfg": ("c3_430k_976x544", (0.4, 2000.0, 10.0), None),
    # depth-complexity stress: 186 M list entries, lists of several thousand per tile (a band of tile
    # rows could be given here to bound the oracle's time: the cotangent is then zero outside it, so the
    # pairs outside contribute exact zeros; the whole 4K frame takes the oracle ~25 s on 16 cores)
    "c5": ("c5_5m_3840x2160", (0.8, 1000.0, 100.0), None),
    # BASELINE config 4's cloud and frame (2M points at 976x544), one view of it
    "c4": ("c4_2m_976x544", (0.8, 1000.0, 100.0), None),
}


@pytest.mark.parametrize("case", list(FULL_SIZE_CASES))
def test_full_size_against_oracle(oracle32, oracle64, case):
    """BASELINE.json configs 2, 3 and 5 at FULL size against the oracle: every forward stage value-exact
    (ids, attributes, tile counts, keys, sorted order, ranges, image, depth, alpha, last index, counts),
    gradients at 1e-4 against the fp64 arbiter (whole tensor, per column and per row), plus
    size-independent properties."""
    import time
    import bench
    workload, (near, far, scale), band = FULL_SIZE_CASES[case]
    n, w, h, sf, seed = bench.WORKLOADS[workload]
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    got = util.cuda_forward_stages(scene, near, far, scale)
    keys = got["point_in_camera_sort_key"]
    assert np.all(keys[1:] >= keys[:-1])
    assert int(got["num_overlap_tiles"].sum()) == keys.shape[0]
    pay = got["point_offset_with_sort_key"]
    assert pay.min() >= 0 and pay.max() < got["point_id_in_camera_list"].shape[0]
    assert np.array_equal(np.bincount(pay, minlength=got["num_overlap_tiles"].shape[0]),
                          got["num_overlap_tiles"])
    assert np.isfinite(got["rasterized_image"]).all()
    lists = (got["tile_points_end"] - got["tile_points_start"]).reshape(h // 16, w // 16)
    per_pixel_len = np.repeat(np.repeat(lists, 16, 0), 16, 1)
    assert np.all(got["pixel_valid_point_count"] <= per_pixel_len)
    t0 = time.time()
    ref = util.oracle_forward(oracle32, scene, near, far, scale)
    t_fwd = time.time() - t0
    report = util.mismatch_report(ref, got)
    assert not report, "\n".join(report)
    m, k = ref["point_id_in_camera_list"].shape[0], keys.shape[0]
    del got

    g = np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32)
    if band is not None:
        g[:16 * band[0]] = 0
        g[16 * band[1]:] = 0
        oracle64.set_tile_row_band(*band)
    ref64 = {key: (v.astype(np.float64) if isinstance(v, np.ndarray) and v.dtype == np.float32 else v)
             for key, v in ref.items()}
    t0 = time.time()
    try:
        b64 = oracle64.backward(ref64, g.astype(np.float64), 3)
    finally:
        oracle64.set_tile_row_band(-1, -1)
    t_bwd = time.time() - t0
    del ref64
    t = util.scene_tensors(scene, requires_grad=True)
    hooked = {}
    ras = _module(near, far, scale, hook=(lambda hk: hooked.update(h=hk)) if band is None else None)
    image, _, _ = _call(ras, scene, t, band=3)
    (image * torch.from_numpy(g).cuda()).sum().backward()
    print(f"{case}: M={m} K={k} pairs fwd {ref['pairs_fwd']} bwd {b64['pairs_bwd']}; "
          f"oracle forward {t_fwd:.1f} s, fp64 backward {t_bwd:.1f} s")
    for name, cuda, o64 in (("xyz", t["point_cloud"].grad.cpu().numpy(), b64["grad_pointcloud"]),
                            ("features", t["point_cloud_features"].grad.cpu().numpy(),
                             b64["grad_pointcloud_features"])):
        l2, mx = util.grad_error(o64, cuda)
        rows = util.row_error_summary(o64, cuda)
        print(f"{case} grad {name}: rel-l2 {l2:.2e} max-col-norm {mx:.2e} per-row p50 {rows['p50']:.1e} "
              f"p99 {rows['p99']:.1e} p99.9 {rows['p999']:.1e} max {rows['max']:.1e} ({rows['rows']} rows)")
        assert l2 <= 1e-4 and mx <= 1e-4
        assert rows["p99"] <= 1e-4
    if band is None:  # the whole frame went through both: the same pairs took part
        assert np.array_equal(hooked["h"].num_affected_pixels.cpu().numpy(), b64["num_affected_pixels"])


def test_accumulate_gradients_into_sums_views():
    """Extension for multi-view batches: backward adds into caller buffers; equals the sum of the
    per-view gradients autograd would have produced."""
    scene = make_scene(5000, 128, 96, seed=21, scale_factor=0.03)
    g = torch.from_numpy(np.random.default_rng(2).uniform(0, 1, (96, 128, 3)).astype(np.float32)).cuda()
    poses = [(0.0, 0.0), (0.02, -0.01), (-0.03, 0.02)]
    ras = _module()

    def run(t, dx, dy):
        t["t_pointcloud_camera"] = torch.tensor([[dx, dy, 0.0]], device="cuda")
        image, _, _ = _call(ras, scene, t, band=3)
        (image * g).sum().backward()

    ref = util.scene_tensors(scene, requires_grad=True)
    for dx, dy in poses:                       # plain autograd accumulation
        run(ref, dx, dy)
    t = util.scene_tensors(scene, requires_grad=True)
    n = t["point_cloud"].shape[0]
    gx = torch.zeros(n, 3, device="cuda")
    gf = torch.zeros(n, 56, device="cuda")
    ras.accumulate_gradients_into(gx, gf)
    for dx, dy in poses:
        run(t, dx, dy)
    ras.accumulate_gradients_into(None, None)
    assert t["point_cloud"].grad is None and t["point_cloud_features"].grad is None
    for got, want in ((gx, ref["point_cloud"].grad), (gf, ref["point_cloud_features"].grad)):
        l2, mx = util.grad_error(want.cpu().numpy(), got.cpu().numpy())
        assert l2 <= 1e-5 and mx <= 1e-5, (l2, mx)
    with pytest.raises(ValueError):
        ras.accumulate_gradients_into(gx[:, :2], gf)


@pytest.mark.parametrize("n,objects,band", [(5000, 1, 3), (7013, 3, 1)])
def test_gradient_batch_equals_per_view_accumulation(n, objects, band):
    """begin_gradient_batch / end_gradient_batch (gs_point_bwd_batch: the per-point chain rule of all
    parked views in one pass over the points, rows stored once) against accumulate_gradients_into
    (gs_point_bwd per view, rows added): the same sums up to the order of the additions -- for one and for
    several point ranges, with a view that sees nothing, several objects, and poisoned output buffers
    (every row of the range must be written, also those of points no view sees)."""
    scene = make_scene(n, 128, 96, seed=21, scale_factor=0.03, num_objects=objects)
    g = torch.from_numpy(np.random.default_rng(2).uniform(0, 1, (96, 128, 3)).astype(np.float32)).cuda()
    poses = [(0.0, 0.0, 0.0), (0.02, -0.01, 0.0), (-0.03, 0.02, 0.1), (0.0, 0.0, 500.0)]   # the last sees nothing: the camera sits beyond every point
    ras = _module()

    def run(t, pose):
        t["t_pointcloud_camera"] = t["t_base"] + torch.tensor([list(pose)], device="cuda")
        image, _, _ = _call(ras, scene, t, band=band)
        (image * g).sum().backward()

    def fresh():
        t = util.scene_tensors(scene, requires_grad=True)
        t["t_base"] = t["t_pointcloud_camera"].clone()
        return t

    ref = fresh()
    gx0 = torch.zeros(n, 3, device="cuda")
    gf0 = torch.zeros(n, 56, device="cuda")
    ras.accumulate_gradients_into(gx0, gf0)
    for pose in poses:
        run(ref, pose)
    ras.accumulate_gradients_into(None, None)
    assert ras.last_forward_info["num_points_in_camera"] == 0
    for chunks in (1, 3):
        t = fresh()
        gx = torch.full((n, 3), float("nan"), device="cuda")
        gf = torch.full((n, 56), float("nan"), device="cuda")
        ras.begin_gradient_batch(gx, gf)
        for pose in poses:
            run(t, pose)
        assert t["point_cloud"].grad is None and t["point_cloud_features"].grad is None
        ranges = []
        assert ras.end_gradient_batch(chunks, lambda i, first, count: ranges.append((i, first, count))) == len(poses)
        torch.cuda.synchronize()
        assert [r[0] for r in ranges] == list(range(len(ranges))) and len(ranges) == chunks
        assert ranges[0][1] == 0 and sum(r[2] for r in ranges) == n
        assert all(a[1] + a[2] == b[1] for a, b in zip(ranges, ranges[1:]))
        assert torch.isfinite(gx).all() and torch.isfinite(gf).all()
        for got, want in ((gx, gx0), (gf, gf0)):
            l2, mx = util.grad_error(want.cpu().numpy(), got.cpu().numpy())
            assert l2 <= 1e-5 and mx <= 1e-5, (chunks, l2, mx)
        keep = {0: 1, 1: 4, 2: 9, 3: 16}[band]
        for base in (8, 24, 40):
            assert bool((gf[:, base + keep:base + 16] == 0).all())
    # no view at all: the gradient of the batch is zero, and it is written
    gx = torch.full((n, 3), float("nan"), device="cuda")
    gf = torch.full((n, 56), float("nan"), device="cuda")
    ras.begin_gradient_batch(gx, gf)
    assert ras.end_gradient_batch() == 0
    assert not gx.any() and not gf.any()
    with pytest.raises(RuntimeError):
        ras.end_gradient_batch()
    with pytest.raises(ValueError, match="hook"):
        _module(hook=lambda h: None).begin_gradient_batch(gx, gf)


@pytest.mark.parametrize("n,objects,band", [(5003, 1, 3), (20000, 3, 2)])
def test_plain_backward_storing_all_rows_equals_zero_fill_plus_scatter(n, objects, band, monkeypatch):
    """A plain backward (no hook: what the unchanged trainer runs) stores every gradient row with
    gs_point_bwd_all_points; GS_STORE_ALL_ROWS=0 keeps the reference's shape of the step (zero fill of the
    N x 59 floats, R.py:1051-1058, then the visible rows).  Same numbers; rows of points the view does not
    see are exact zeros either way."""
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    scene = make_scene(n, 160, 128, seed=41, scale_factor=0.03, num_objects=objects)
    g = torch.from_numpy(np.random.default_rng(6).uniform(0, 1, (128, 160, 3)).astype(np.float32)).cuda()
    grads = {}
    for store_all in (False, True):
        monkeypatch.setattr(impl, "STORE_ALL_ROWS", store_all)
        ras = _module()
        t = util.scene_tensors(scene, requires_grad=True)
        image, _, _ = _call(ras, scene, t, band=band)
        (image * g).sum().backward()
        torch.cuda.synchronize()
        grads[store_all] = (t["point_cloud"].grad.clone(), t["point_cloud_features"].grad.clone())
        assert ras.last_forward_info["num_points_in_camera"] < n
    for got, want in zip(grads[True], grads[False]):
        assert torch.isfinite(got).all()
        l2, mx = util.grad_error(want.cpu().numpy(), got.cpu().numpy())
        assert l2 <= 1e-6 and mx <= 1e-6, (l2, mx)
        unseen = (want == 0).all(dim=1)
        assert bool(unseen.any()) and bool((got[unseen] == 0).all())


def test_view_sharded_step_on_two_lanes_matches_one_lane():
    """distributed.ViewShardedTrainerStep with the fused accumulation: six views spread over two CUDA
    streams give the gradient of the one-stream run (up to the order of the RED adds)."""
    from taichi_3d_gaussian_splatting_b200.distributed import ViewShardedTrainerStep
    scene = make_scene(20000, 256, 160, seed=23, scale_factor=0.02)
    g = torch.from_numpy(np.random.default_rng(4).uniform(0, 1, (160, 256, 3)).astype(np.float32)).cuda()
    views = [(0.01 * i, -0.005 * i) for i in range(6)]
    results = []
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    for lanes, priority_sections in ((1, False), (2, False), (3, False), (3, True), (1, True)):
        ras = _module()
        t = util.scene_tensors(scene, requires_grad=True)
        # (3, True): the small kernels of every view on a high-priority companion stream of its lane
        impl.PRIORITY_SECTIONS = priority_sections

        def render(xyz, feat, view, t=t, ras=ras):
            t["t_pointcloud_camera"] = torch.tensor([[view[0], view[1], 0.0]], device="cuda")
            image, _, _ = _call(ras, scene, t, band=3)
            return image

        step = ViewShardedTrainerStep(render, lambda image, view: (image * g).sum(), rasteriser=ras, lanes=lanes)
        res = step(t["point_cloud"], t["point_cloud_features"], views)
        torch.cuda.synchronize()
        assert res.num_local_views == 6
        impl.PRIORITY_SECTIONS = False
        results.append((t["point_cloud"].grad.clone(), t["point_cloud_features"].grad.clone(), float(res.loss)))
    for gx, gf, loss in results[1:]:
        assert abs(loss - results[0][2]) <= 1e-5 * abs(results[0][2])
        for got, want in ((gx, results[0][0]), (gf, results[0][1])):
            l2, mx = util.grad_error(want.cpu().numpy(), got.cpu().numpy())
            assert l2 <= 1e-5 and mx <= 1e-5, (l2, mx)


@pytest.mark.parametrize("n", [5001, 5002, 5003])
def test_fused_step_with_point_counts_that_are_not_multiples_of_four(n):
    """The flat N x 59 buffer of the fused step holds the N x 56 block first, so its rows stay 16-byte
    aligned for every N (they are added with red.global.add.v4); with the N x 3 block first any N that is
    not a multiple of 4 faulted."""
    from taichi_3d_gaussian_splatting_b200.distributed import ViewShardedTrainerStep
    scene = make_scene(n, 128, 96, seed=31, scale_factor=0.03)
    g = torch.from_numpy(np.random.default_rng(2).uniform(0, 1, (96, 128, 3)).astype(np.float32)).cuda()
    views = [(0.0, 0.0), (0.02, -0.01), (-0.03, 0.02)]
    grads = []
    for fused in (False, True):
        ras = _module()
        t = util.scene_tensors(scene, requires_grad=True)

        def render(xyz, feat, view, t=t, ras=ras):
            t["t_pointcloud_camera"] = torch.tensor([[view[0], view[1], 0.0]], device="cuda")
            return _call(ras, scene, t, band=3)[0]

        step = ViewShardedTrainerStep(render, lambda image, view: (image * g).sum(),
                                      rasteriser=ras if fused else None, lanes=2)
        step(t["point_cloud"], t["point_cloud_features"], views)
        torch.cuda.synchronize()
        if fused:
            assert t["point_cloud_features"].grad.data_ptr() % 16 == 0
        grads.append((t["point_cloud"].grad.cpu().numpy(), t["point_cloud_features"].grad.cpu().numpy()))
    for a, b in zip(grads[0], grads[1]):
        l2, mx = util.grad_error(a, b)
        assert l2 <= 1e-5 and mx <= 1e-5, (l2, mx)
    with pytest.raises(ValueError, match="16-byte"):
        ras = _module()
        flat = torch.zeros(n * 59 + 1, device="cuda")
        ras.accumulate_gradients_into(flat[1 + 56 * n:].view(n, 3), flat[1:1 + 56 * n].view(n, 56))


def test_fused_step_keeps_gradients_from_outside_the_rasteriser():
    """A loss term on the parameters that does not pass through the rasteriser (the reference trainer's
    scale regulariser, LossFunction.py:40-52) must survive the fused step's in-place accumulation."""
    from taichi_3d_gaussian_splatting_b200.distributed import ViewShardedTrainerStep
    scene = make_scene(4000, 128, 96, seed=33, scale_factor=0.03)
    g = torch.from_numpy(np.random.default_rng(3).uniform(0, 1, (96, 128, 3)).astype(np.float32)).cuda()
    views = [(0.0, 0.0), (0.02, -0.01)]
    grads = []
    for fused in (False, True):
        ras = _module()
        t = util.scene_tensors(scene, requires_grad=True)

        def render(xyz, feat, view, t=t, ras=ras):
            t["t_pointcloud_camera"] = torch.tensor([[view[0], view[1], 0.0]], device="cuda")
            return _call(ras, scene, t, band=3)[0]

        def loss_fn(image, view, t=t):
            reg = torch.exp(t["point_cloud_features"][:, 4:7]).norm(dim=1).mean() + t["point_cloud"].pow(2).mean()
            return (image * g).sum() + 2.0 * reg

        step = ViewShardedTrainerStep(render, loss_fn, rasteriser=ras if fused else None, lanes=1)
        step(t["point_cloud"], t["point_cloud_features"], views)
        grads.append((t["point_cloud"].grad.cpu().numpy(), t["point_cloud_features"].grad.cpu().numpy()))
    assert np.abs(grads[0][1][:, 4:7]).max() > 0
    for a, b in zip(grads[0], grads[1]):
        l2, mx = util.grad_error(a, b)
        assert l2 <= 1e-5 and mx <= 1e-5, (l2, mx)


def test_densification_statistics_with_views_on_concurrent_lanes():
    """The backward hook of every view runs on its view's lane; AdaptiveControllerAccumulators.update adds
    with atomics, so views whose hooks run concurrently lose no update (integer statistics exactly equal,
    float sums equal up to the order of the additions)."""
    from taichi_3d_gaussian_splatting_b200.distributed import ViewShardedTrainerStep
    from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators
    scene = make_scene(30000, 256, 160, seed=35, scale_factor=0.02)
    g = torch.from_numpy(np.random.default_rng(4).uniform(0, 1, (160, 256, 3)).astype(np.float32)).cuda()
    views = [(0.004 * i, -0.003 * i) for i in range(8)]
    stats = []
    for lanes in (1, 4):
        t = util.scene_tensors(scene, requires_grad=True)
        acc = AdaptiveControllerAccumulators(t["point_cloud"].shape[0], "cuda")
        ras = _module(hook=acc.update)

        def render(xyz, feat, view, t=t, ras=ras):
            t["t_pointcloud_camera"] = torch.tensor([[view[0], view[1], 0.0]], device="cuda")
            return _call(ras, scene, t, band=3)[0]

        step = ViewShardedTrainerStep(render, lambda image, view: (image * g).sum(), rasteriser=ras, lanes=lanes)
        for _ in range(3):
            step(t["point_cloud"], t["point_cloud_features"], views)
        torch.cuda.synchronize()
        stats.append({k: v.cpu().numpy() for k, v in vars(acc).items() if isinstance(v, torch.Tensor)})
    assert len(stats[0]) >= 6
    for key, a in stats[0].items():
        b = stats[1][key]
        if np.issubdtype(a.dtype, np.integer):
            assert np.array_equal(a, b), key
        else:
            assert np.allclose(a, b, rtol=1e-4, atol=1e-7 * max(1.0, np.abs(a).max())), key
    assert stats[0]["accumulated_num_in_camera"].max() == 24


def test_persistent_blend_launch_equals_one_cta_per_tile(monkeypatch):
    """The backward blend's persistent launch (warps pull (tile, part) units from a per-stream counter
    that the last warp re-arms) against the plain one-CTA-per-tile launch (GS_BLEND_PERSISTENT=0):
    gradients equal up to the order of the RED adds, the same pairs visited; repeated launches on one
    stream and an image with fewer tiles than resident CTAs included."""
    for (n, w, h, sf) in ((30000, 256, 160, 0.015), (500, 32, 16, 0.05), (200000, 1024, 512, 0.006)):
        scene = make_scene(n, w, h, seed=41, scale_factor=sf)
        g = torch.from_numpy(np.random.default_rng(8).uniform(0, 1, (h, w, 3)).astype(np.float32)).cuda()
        runs = []
        for mode in ("1", "0", "1"):
            monkeypatch.setenv("GS_BLEND_PERSISTENT", mode)
            hooked = {}
            ras = _module(hook=lambda hk: hooked.update(h=hk))
            t = util.scene_tensors(scene, requires_grad=True)
            image, depth, count = _call(ras, scene, t, band=3)
            (image * g).sum().backward()
            torch.cuda.synchronize()
            runs.append((image.detach().cpu().numpy(), depth.detach().cpu().numpy(), count.cpu().numpy(),
                         t["point_cloud"].grad.cpu().numpy(), t["point_cloud_features"].grad.cpu().numpy(),
                         hooked["h"].num_affected_pixels.cpu().numpy()))
        for other in runs[1:]:
            for a, b in zip(runs[0][:3], other[:3]):
                assert np.array_equal(a, b)
            assert np.array_equal(runs[0][5], other[5])
            for a, b in zip(runs[0][3:5], other[3:5]):
                l2, mx = util.grad_error(a, b)
                assert l2 <= 1e-5 and mx <= 1e-5, (l2, mx)


@pytest.mark.parametrize("workload", ["c4_2m_976x544", "c5_5m_3840x2160"])
def test_full_size_properties_of_the_other_baseline_configs(workload):
    """BASELINE.json configs 3-5 at full size, through size-independent properties (the oracle
    comparison at these sizes is config 2's, above): the sorted keys are sorted and stable, the tile
    ranges partition them, every point's tile count is accounted for, per-pixel counts are bounded
    by the tile list, the render is finite and deterministic, and a backward pass touches exactly
    the visible points."""
    import bench
    n, w, h, sf, seed = bench.WORKLOADS[workload]
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    t = util.scene_tensors(scene, requires_grad=True)
    ras = _module()
    image, depth, count = _call(ras, scene, t, band=3)
    state_info = ras.last_forward_info
    m, k = state_info["num_points_in_camera"], state_info["num_keys"]
    assert 0 < m <= n and k > 0
    image.sum().backward()
    torch.cuda.synchronize()
    img = image.detach()
    assert torch.isfinite(img).all() and torch.isfinite(depth).all()
    assert int(count.min()) >= 0
    gx, gf = t["point_cloud"].grad, t["point_cloud_features"].grad
    assert torch.isfinite(gx).all() and torch.isfinite(gf).all()
    touched = int((gf.abs().sum(dim=1) > 0).sum())
    assert 0 < touched <= m                      # only visible points receive a gradient
    assert int((gf[t["point_invalid_mask"] != 0].abs().sum() > 0)) == 0
    # the same forward again from the same inputs (a fresh copy: the first call normalised the
    # quaternion columns in place): bit-identical outputs, the forward has no atomics
    with torch.no_grad():
        image2, depth2, count2 = _call(ras, scene, util.scene_tensors(scene), band=3)
    assert torch.equal(image2, img) and torch.equal(depth2, depth) and torch.equal(count2, count)
    # stage-level invariants on the device buffers
    st = util.cuda_forward_state(scene) if hasattr(util, "cuda_forward_state") else None
    if st is not None:
        keys = st.sorted_keys()
        assert bool((keys[1:] >= keys[:-1]).all())
        lens = (st.tile_end - st.tile_start).to(torch.int64)
        assert int(lens.sum()) == st.k
        per_pixel = lens.view(h // 16, w // 16).repeat_interleave(16, 0).repeat_interleave(16, 1)
        assert bool((st.count.to(torch.int64) <= per_pixel).all())
        tiles_per_point = torch.bincount(st.payload.to(torch.int64), minlength=st.m)
        assert torch.equal(tiles_per_point.to(torch.int32), st.buffers.view("num_tiles", st.m).view(-1))


def _blend_case_by_hand(seed=0):
    """A 32x32 image (4 tiles) whose tiles list (nearly) every splat, with three kinds of entries the
    synthetic scenes never produce: (a) conics that are NOT positive definite (an exponent above zero:
    only the guarded literal path may see it), (b) a NaN opacity (the forward takes such a pair --
    `alpha < 1/255` is false -- the backward skips it -- `alpha >= 1/255` is false: R.py:451 vs :634),
    (c) splats whose alpha sits within an ulp of 1/255 on a ring of pixel centres."""
    rng = np.random.default_rng(seed)
    W = H = 32
    m_regular, m_edge = 30, 120
    uv, conic, alpha, color, z = [], [], [], [], []
    for _ in range(m_regular):
        uv.append(rng.uniform(-2, 34, 2))
        a, c = rng.uniform(0.004, 0.08, 2)
        b = rng.uniform(-0.9, 0.9) * np.sqrt(a * c)
        conic.append([a, b, c, rng.uniform(0.6, 1.0)])
        alpha.append(rng.uniform(0.02, 0.3))
    for _ in range(m_edge):  # isotropic, centred on a pixel centre: alpha == 1/255 (+- rounding) at distance d
        centre = rng.integers(4, 28, 2) + 0.5
        a = rng.uniform(0.01, 0.06)
        d2 = float(rng.choice([1.0, 2.0, 4.0, 5.0, 8.0, 9.0, 10.0, 13.0]))
        rescale = rng.uniform(0.7, 1.0)
        p = np.exp(-0.5 * a * d2) * rescale
        uv.append(centre)
        conic.append([a, 0.0, a, rescale])
        alpha.append(min((1.0 / 255.0) / p * (1.0 + rng.uniform(-2e-7, 2e-7)), 0.95))
    irregular = [
        ([10.5, 12.5], [-0.002, 0.0, 0.05, 0.9], 0.5),      # saddle: exponent > 0 away from the centre row
        ([20.2, 7.7], [0.03, 0.06, 0.03, 0.8], 0.4),        # b^2 > a c
        ([5.0, 25.0], [-0.001, 0.0005, -0.002, 1.0], 0.3),  # negative definite: alpha grows outwards
        ([16.0, 16.0], [0.02, 0.0, 0.02, 0.9], np.nan),     # NaN opacity
    ]
    for c_uv, c_conic, c_alpha in irregular:
        uv.append(c_uv); conic.append(c_conic); alpha.append(c_alpha)
    m = len(uv)
    uv = np.asarray(uv, np.float32)
    conic = np.asarray(conic, np.float32)
    alpha = np.asarray(alpha, np.float32)
    color = rng.uniform(0.05, 1.0, (m, 3)).astype(np.float32)
    xyz_cam = np.zeros((m, 3), np.float32)
    xyz_cam[:, 2] = rng.uniform(1, 20, m).astype(np.float32)
    # tiles 0 and 1 list every splat, tiles 2 and 3 only the regular and edge ones: the backward must take
    # its whole-quarter literal path in the former and its per-pixel one (edge pairs) in the latter
    order = rng.permutation(m).astype(np.int32)
    regular_only = order[order < m - len(irregular)]
    lists = [order, order, regular_only, regular_only]
    payload = np.concatenate(lists).astype(np.int32)
    lens = np.array([len(x) for x in lists])
    end = np.cumsum(lens).astype(np.int32)
    start = (end - lens).astype(np.int32)
    records = np.zeros((m, 12), np.float32)
    records[:, 0:2] = uv
    records[:, 2:5] = conic[:, 0:3]
    records[:, 5] = conic[:, 3]
    records[:, 6] = alpha
    records[:, 7:10] = color
    records[:, 10] = xyz_cam[:, 2]
    records[:, 11] = util.decision_band_log2(conic, alpha)
    return dict(W=W, H=H, m=m, uv=uv, conic=conic, alpha=alpha, color=color, xyz_cam=xyz_cam,
                payload=payload, start=start, end=end, records=records, first_irregular=m_regular + m_edge)


def test_persistent_backward_recovers_from_a_launch_that_left_its_work_counter_half_way(monkeypatch):
    """The persistent backward blend pulls its units from a {next, done} pair that the last warp of a
    launch re-arms (csrc/gs_common.cuh).  A launch that dies before that leaves the pair at arbitrary
    values; the launcher records the failure and the next call on that stream zeroes the pair first.
    GS_TEST_POISON_WORK_COUNTER=1 makes one call behave like such a launch (pair overwritten, error
    returned): without the re-arm the following call would find next >= units, do nothing and return
    zeros."""
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    capi = _lib()
    lib = capi.lib
    monkeypatch.setenv("GS_BLEND_PERSISTENT", "1")
    scene = make_scene(20000, 320, 240, seed=77, scale_factor=0.01)
    t = util.scene_tensors(scene)
    cfg = impl.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig()
    cam = CameraInfo(camera_intrinsics=t["camera_intrinsics"], camera_height=240, camera_width=320, camera_id=0)
    st = impl._forward_impl(cfg, t["point_cloud"], t["point_cloud_features"], t["point_invalid_mask"],
                            t["point_object_id"], t["q_pointcloud_camera"], t["t_pointcloud_camera"], cam,
                            keep_cell_masks=True)
    g = torch.from_numpy(np.random.default_rng(5).uniform(0, 1, (240, 320, 3)).astype(np.float32)).cuda()
    scratch = torch.empty(int(lib.gs_blend_bwd_scratch_bytes(240, 320)), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def bwd():
        grads = torch.zeros(st.m, 12, dtype=torch.float32, device="cuda")
        status = lib.gs_blend_bwd(240, 320, st.tile_start.data_ptr(), st.tile_end.data_ptr(),
                                  st.payload.data_ptr(), st.buffers.ptr("records"), g.data_ptr(),
                                  st.acc_alpha.data_ptr(), st.last.data_ptr(), grads.data_ptr(), None,
                                  st.cell_masks.data_ptr(), st.k, st.quarter_flags.data_ptr(),
                                  scratch.data_ptr(), stream)
        torch.cuda.synchronize()
        return status, grads

    status, want = bwd()
    assert status == 0 and float(want.abs().sum()) > 0
    monkeypatch.setenv("GS_TEST_POISON_WORK_COUNTER", "1")
    status, nothing = bwd()
    assert status != 0 and not nothing.any()
    monkeypatch.delenv("GS_TEST_POISON_WORK_COUNTER")
    for _ in range(2):   # the call after the failure, and one more (the pair re-arms itself again)
        status, got = bwd()
        assert status == 0
        l2, mx = util.grad_error(want.cpu().numpy(), got.cpu().numpy())
        assert l2 <= 1e-6 and mx <= 1e-6, (l2, mx)


@pytest.mark.parametrize("magnitude", [True, False])
@pytest.mark.parametrize("persistent", ["1", "0"])
def test_blend_backward_decides_like_the_reference_on_irregular_and_edge_entries(
        oracle32, oracle64, magnitude, persistent, monkeypatch):
    """gs_blend_fwd + gs_blend_bwd through the C ABI on hand-made records: irregular conics take the
    guarded literal path in BOTH directions and pairs on the alpha = 1/255 edge are re-decided with
    the reference's backward arithmetic (R.py:626-669, utils.py:331-345), so num_affected_pixels
    equals the literal oracle's and the gradients match the fp64 arbiter."""
    capi = _lib()
    monkeypatch.setenv("GS_BLEND_PERSISTENT", persistent)
    c = _blend_case_by_hand()
    W, H, m = c["W"], c["H"], c["m"]
    dev = torch.device("cuda")
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    start, end, payload, records = tt(c["start"]), tt(c["end"]), tt(c["payload"]), tt(c["records"])
    image = torch.empty(H, W, 3, device=dev)
    depth, acc = torch.empty(H, W, device=dev), torch.empty(H, W, device=dev)
    last = torch.empty(H, W, dtype=torch.int32, device=dev)
    count = torch.empty(H, W, dtype=torch.int32, device=dev)
    k = payload.shape[0]
    masks = torch.zeros(4 * k, dtype=torch.uint8, device=dev)
    flags = torch.full((4 * 4,), 7, dtype=torch.uint8, device=dev)       # every byte must be written
    scratch = torch.full((int(capi.lib.gs_blend_bwd_scratch_bytes(H, W)),), 3, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    capi.check(capi.lib.gs_blend_fwd(H, W, start.data_ptr(), end.data_ptr(), payload.data_ptr(),
                                     records.data_ptr(), image.data_ptr(), depth.data_ptr(), acc.data_ptr(),
                                     last.data_ptr(), count.data_ptr(), 0, masks.data_ptr(), k,
                                     flags.data_ptr(), st), "gs_blend_fwd")
    fl = flags.cpu().numpy().reshape(4, 4)
    assert set(fl.ravel().tolist()) <= {0, 1} and fl[:2].all() and not fl[2:].any()
    f = oracle32.blend_fwd(H, W, c["start"], c["end"], c["payload"], c["uv"], c["xyz_cam"], c["conic"],
                           c["alpha"], c["color"])
    got = dict(rasterized_image=image.cpu().numpy(), rasterized_depth=depth.cpu().numpy(),
               pixel_accumulated_alpha=acc.cpu().numpy(),
               pixel_offset_of_last_effective_point=last.cpu().numpy(),
               pixel_valid_point_count=count.cpu().numpy())
    report = util.mismatch_report(f, got, keys=list(got))
    assert not report, "\n".join(report)

    g = np.random.default_rng(9).uniform(0.1, 1, (H, W, 3)).astype(np.float32)
    ids = np.arange(m, dtype=np.int32)
    b32 = oracle32.blend_bwd(H, W, m, c["start"], c["end"], c["payload"], ids, g,
                             f["pixel_accumulated_alpha"], f["pixel_offset_of_last_effective_point"],
                             c["uv"], c["conic"], c["alpha"], c["color"])
    d = lambda a: a.astype(np.float64)
    b64 = oracle64.blend_bwd(H, W, m, c["start"], c["end"], c["payload"], ids, d(g),
                             d(f["pixel_accumulated_alpha"]), f["pixel_offset_of_last_effective_point"],
                             d(c["uv"]), d(c["conic"]), d(c["alpha"]), d(c["color"]))
    assert np.array_equal(b32["num_affected_pixels"], b64["num_affected_pixels"])
    splat_grads = torch.zeros(m, 12, device=dev)
    on_image = torch.empty(H, W, 2, device=dev)
    capi.check(capi.lib.gs_blend_bwd(H, W, start.data_ptr(), end.data_ptr(), payload.data_ptr(),
                                     records.data_ptr(), tt(g).data_ptr(), acc.data_ptr(), last.data_ptr(),
                                     splat_grads.data_ptr(), on_image.data_ptr() if magnitude else None,
                                     masks.data_ptr(), k, flags.data_ptr(), scratch.data_ptr(), st),
               "gs_blend_bwd")
    torch.cuda.synchronize()
    sg = splat_grads.cpu().numpy()
    aff = sg[:, 10].astype(np.int64)
    assert np.array_equal(aff, b32["num_affected_pixels"]), np.nonzero(aff != b32["num_affected_pixels"])
    # the irregular entries did take part (otherwise this test checks nothing); the NaN-opacity one
    # is skipped by the backward, as in the reference
    first = c["first_irregular"]
    assert (aff[first:first + 3] > 0).all() and aff[first + 3] == 0
    assert np.isfinite(sg).all()
    want = np.concatenate([b64["grad_uv"], b64["cov_buf"], b64["color_buf"], b64["grad_logit"][:, None],
                           b64["magnitude_grad_viewspace"][:, None]], axis=1)
    cols = 10 if magnitude else 9
    l2, mx = util.grad_error(want[:, :cols], sg[:, :cols])
    rows = util.row_error_summary(want[:, :cols], sg[:, :cols])
    print(f"hand-made blend case: rel-l2 {l2:.2e} max-col-norm {mx:.2e} per-row p99.9 {rows['p999']:.1e}")
    assert l2 <= 1e-4 and mx <= 1e-4
    if magnitude:
        l2, mx = util.grad_error(b64["magnitude_grad_viewspace_on_image"].reshape(-1, 2),
                                 on_image.cpu().numpy().reshape(-1, 2))
        assert l2 <= 1e-4 and mx <= 1e-4
    else:
        assert np.all(sg[:, 9] == 0)
    # the scratch holds the list of pixels the literal pass re-did ({count, entries}; bit 31 = a pixel of
    # a quarter left whole to it): this case must have exercised both kinds of literal work
    fix = scratch.cpu().numpy().view(np.uint32)
    entries = fix[1:1 + int(fix[0])]
    whole = entries[entries >> 31 == 1] & 0x7fffffff
    edge = entries[entries >> 31 == 0]
    assert len(np.unique(entries & 0x7fffffff)) == len(entries)          # every pixel at most once
    assert len(whole) == 2 * 256 and set((whole // W // 16 * 2 + whole % W // 16).tolist()) == {0, 1}
    assert len(edge) > 20 and set((edge // W // 16 * 2 + edge % W // 16).tolist()) <= {2, 3}
