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
