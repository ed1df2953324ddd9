"""Ingest: the index rule of PIL's NEAREST resize (dataset_mapper.py:127-128) on CPU, and the device
gather against PIL on the GPU."""
import numpy as np
import pytest
import torch
from PIL import Image

from tdod import ingest

SHAPES = [(480, 640, 530, 730), (530, 730, 480, 640), (375, 1242, 384, 1280), (1080, 1920, 530, 730), (100, 37, 311, 53),
          (7, 9, 640, 480), (427, 561, 512, 672)]


@pytest.mark.parametrize("h,w,H,W", SHAPES)
def test_index_maps_match_pil(h, w, H, W):
    a = np.arange(h * w, dtype=np.float32).reshape(h, w)          # exact in fp32 below 2^24
    r = np.array(Image.fromarray(a).resize((W, H), Image.NEAREST))
    assert np.array_equal(a[ingest.pil_nearest_map(h, H)][:, ingest.pil_nearest_map(w, W)], r)


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,H,W", SHAPES[:5])
def test_device_resize_matches_pil(h, w, H, W):
    rs = np.random.RandomState(h + w)
    depth = rs.uniform(0.3, 20, (3, h, w)).astype(np.float32)
    ground = rs.rand(3, h, w) > 0.6
    d = torch.device("cuda:0")
    got_d = ingest.resize_nearest(torch.from_numpy(depth).to(d), H, W).cpu().numpy()
    got_g = ingest.resize_nearest(torch.from_numpy(ground).to(d), H, W).cpu().numpy()
    for k in range(3):
        assert np.array_equal(got_d[k], np.array(Image.fromarray(depth[k]).resize((W, H), Image.NEAREST)))
        assert np.array_equal(got_g[k], np.array(Image.fromarray(ground[k]).resize((W, H), Image.NEAREST)))
    one = ingest.resize_nearest(torch.from_numpy(depth[0]).to(d), H, W)
    assert one.shape == (H, W) and np.array_equal(one.cpu().numpy(), got_d[0])


def _write_maps(tmp_path, n, h, w, seed=0):
    rs = np.random.RandomState(seed)
    items, arrays = [], []
    for k in range(n):
        depth = rs.uniform(0.3, 20, (h, w)).astype(np.float32)
        mask = rs.rand(h, w) > 0.6
        dp, gp = tmp_path / ("d%d.npz" % k), tmp_path / ("g%d.npz" % k)
        np.savez_compressed(dp, depth=depth)                 # generate_depth_maps.py:86
        np.savez_compressed(gp, mask=mask)                   # generate_ground_segmentations.py:200
        items.append((str(dp), str(gp) if k % 3 else None))  # some frames have no ground map (dataset_mapper.py:146)
        arrays.append((depth, mask if k % 3 else None))
    return items, arrays


def test_load_maps_reads_the_reference_file_format(tmp_path):
    items, arrays = _write_maps(tmp_path, 4, 37, 53)
    for (dp, gp), (depth, mask) in zip(items, arrays):
        d, g = ingest.load_maps(dp, gp)
        assert d.dtype == np.float32 and np.array_equal(d, depth)
        assert (g is None) == (mask is None)
        if g is not None:
            assert g.dtype == np.uint8 and np.array_equal(g.astype(bool), mask)
    with pytest.raises(ValueError):
        ingest.MapPrefetcher(items, 10, 10, "cpu")


@pytest.mark.gpu
def test_prefetcher_yields_the_dataset_mapper_maps_in_order(tmp_path):
    h, w, H, W = 120, 160, 530, 730
    items, arrays = _write_maps(tmp_path, 11, h, w, seed=3)
    pf = ingest.MapPrefetcher(items, H, W, "cuda:0", workers=3, ahead=3)
    n = 0
    for (d, g), (depth, mask) in zip(pf, arrays):
        # dataset_mapper.py:125-147
        assert np.array_equal(d.cpu().numpy(), np.array(Image.fromarray(depth).resize((W, H), Image.NEAREST)))
        assert (g is None) == (mask is None)
        if g is not None:
            assert np.array_equal(g.cpu().numpy().astype(bool), np.array(Image.fromarray(mask).resize((W, H), Image.NEAREST)))
        n += 1
    assert n == 11 and pf.bytes_h2d == sum(a.nbytes + (0 if m is None else m.nbytes) for a, m in arrays)
