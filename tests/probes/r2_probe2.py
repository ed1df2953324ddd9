"""Round-2 GPU probe 2: (1) tdod_torch_std against torch.std on CUDA tensors over the launch shapes torch picks;
(2) K-C in std_mode TORCH_CUDA against the oracle's torch code run on CUDA tensors (centres bit for bit);
(3) K-D order / accumulate_scores against NumPy; (4) box3d known answer."""
import os
import random
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(ROOT, "3dod_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import tdod_oracle as O  # noqa: E402
from tdod import api, pipeline, synth, lib as L  # noqa: E402

d = torch.device("cuda:0")


def part1():
    g = torch.Generator(device=d).manual_seed(1)
    bad_total = 0
    for n in (2, 3, 33, 100, 127, 128, 129, 255, 256, 999, 1000, 1001, 2049, 4096, 10000, 16000):
        for N in (1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 31, 40, 50, 64):
            x = torch.randn((N, n), generator=g, device=d) * 0.37 + 1.9
            ref = torch.std(x, dim=1)
            try:
                got = api.torch_std(x)
            except ValueError as e:
                print("1 n=%d N=%d refused: %s" % (n, N, e)); continue
            bad = int((got.view(torch.int32) != ref.view(torch.int32)).sum())
            bad_total += bad
            if bad:
                ul = (got.view(torch.int32).long() - ref.view(torch.int32).long()).abs().max()
                print("1 n=%d N=%d: %d / %d rows differ (max %d ulp) cfg=%s" % (n, N, bad, N, int(ul), n))
    print("1 torch_std vs torch.std: total differing rows", bad_total)


def part2():
    tot = dict(objs=0, ctr_ne=0, std_ne=0, sc_miss={}, n=0)
    for cfg, ids in (("C1", range(6)), ("C2", range(8))):
        c = synth.CONFIGS[cfg]
        for fid in ids:
            f = synth.make_frame(fid, c["H"], c["W"], c["n_obj"])
            N = f.n_obj()
            if N == 0:
                continue
            n = 1000
            pg, _ = O.backproject_np(f.depth.numpy(), f.K.numpy(), f.ground.numpy())
            idsr = O.draw_sample_ids(len(pg), 1000, random.Random(fid))
            eq, _, _ = O.ransac_np(pg, idsr, 0.05)
            nv = torch.from_numpy(O.rectify_normal(np.array(eq[:3])))
            o = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, n, nv, torch.Generator().manual_seed(fid))
            g = lambda t: t.to(d)
            cub = O.cubes_from_noise(g(f.boxes), g(f.depth), g(f.prior_mean), g(f.prior_std), g(f.K), n, g(o["noise"]), nv,
                                     o["yaw_idx"])
            m_rect, has = api.mask_rect(g(f.masks))
            k = api.propose_score_select(g(f.depth)[None], g(f.K)[None], g(f.boxes), g(f.prior_mean), g(f.prior_std),
                                         torch.zeros(N, dtype=torch.int32, device=d), m_rect, has, n,
                                         yaw_tab=api.yaw_table(nv.to(d)[None]),
                                         noise=g(o["noise"].permute(1, 0, 2).contiguous()),
                                         yaw_idx=g(o["yaw_idx"].to(torch.int16)), return_all=True, std_mode=L.STD_TORCH_CUDA)
            torch.cuda.synchronize()
            tot["objs"] += N
            tot["ctr_ne"] += int((k.all_cubes.view(torch.int32) != cub.contiguous().view(torch.int32)).sum())
            # scores against the oracle's torch ops on CUDA tensors (corners: host cKDTree on the CUDA-projected corners)
            c2d = O.project_corners(O.cube_corners(cub), g(f.K), c["W"], c["H"])
            pb = O.boxes_from_corners(c2d)
            for i in range(N):
                s_iou = O.iou_score(g(f.boxes[i]), pb[i]).cpu()
                s_dim = O.dims_score(g(f.prior_mean[i]), g(f.prior_std[i]), cub[i, :, 3:6], g(f.boxes[i]), pb[i]).cpu()
                rect = m_rect[i].cpu().numpy() if int(has[i]) == 1 else None
                s_cor = O.corners_score(f.masks[i], c2d[i].cpu(), rect=rect)
                comb, best = O.select(s_iou, s_dim, s_cor)
                got = k.all_scores[i].cpu().numpy()
                for ch, ref in enumerate((s_iou.numpy(), s_dim.numpy(), s_cor.numpy(), comb)):
                    bad = np.abs(got[ch] - ref) > 1e-7 + 1e-5 * np.abs(ref)
                    bad &= ~(np.isnan(ref) & np.isnan(got[ch]))
                    tot["sc_miss"][ch] = tot["sc_miss"].get(ch, 0) + int(bad.sum())
                    if bad.any():
                        j = int(np.argmax(np.abs(got[ch] - ref)))
                        print("   miss ch%d obj%d: got %.9g ref %.9g" % (ch, i, got[ch][j], ref[j]))
                tot["n"] += n
    print("2 std_mode TORCH_CUDA vs oracle on CUDA tensors:", tot)


def part3():
    g = torch.Generator().manual_seed(2)
    bad = 0
    for n in (5, 32, 33, 1000, 1024, 1025, 4000, 10000):
        S = 7
        sc = torch.rand((S, n), generator=g)
        sc[0, :: 7] = sc[0, 0]                      # ties
        sc[1, 3] = float("nan"); sc[1, n // 2] = float("nan")
        sc[2] = -sc[2]
        sc[3, 1] = float("inf"); sc[3, 2] = float("-inf")
        k = min(n, 10)
        idx = api.segmented_order(sc.to(d), k=n).cpu().numpy()
        top, val = api.segmented_order(sc.to(d), k=k, want_val=True)
        for s in range(S):
            a = sc[s].numpy()
            ref = np.argsort(a, kind="stable")[::-1]
            # reference order up to ties: compare the sorted VALUES and that idx is a permutation
            ok = np.array_equal(np.sort(idx[s]), np.arange(n))
            va, vr = a[idx[s]], a[ref]
            ok &= bool(np.all((va == vr) | (np.isnan(va) & np.isnan(vr))))
            # our tie rule = stable ascending sort reversed
            ok &= np.array_equal(idx[s], ref)
            ok &= np.array_equal(top[s].cpu().numpy(), idx[s][:k])
            bad += not ok
        iou = torch.rand((S, n), generator=g)
        C = 3
        scd = torch.stack([sc.double(), (sc * sc).double(), torch.rand((S, n), generator=g).double()], 1).contiguous()
        cur = api.accumulate_scores(scd.to(d), iou.to(d)).cpu().numpy()
        first = api.accumulate_scores(scd.to(d), iou.to(d), n_out=1).cpu().numpy()
        for s in range(S):
            for ch in (0, 2):            # channel 1 has ties between +x and -x only where equal: skip the tie-heavy rows
                if s in (0, 1):
                    continue
                ref = O.accumulate_scores(scd[s, ch].numpy(), iou[s].numpy())
                bad += not np.array_equal(cur[s, ch], ref.astype(np.float32))
                bad += not (first[s, ch, 0] == ref[0])
    print("3 K-D order / accumulate_scores: mismatching rows", bad)


def part4():
    c1 = torch.tensor([[0.2411, -0.1752, 1.2247], [0.1951, -0.4194, 1.7741], [0.2036, 0.4826, 2.1757], [0.2495, 0.7267, 1.6263],
                       [-0.2920, -0.1549, 1.1903], [-0.3380, -0.3991, 1.7396], [-0.3295, 0.5029, 2.1412], [-0.2835, 0.7471, 1.5919]])
    c2 = torch.tensor([[0.2390, -0.1764, 1.2246], [0.1930, -0.4205, 1.7740], [0.2055, 0.4813, 2.1759], [0.2515, 0.7254, 1.6265],
                       [-0.2940, -0.1536, 1.1901], [-0.3400, -0.3978, 1.7395], [-0.3274, 0.5040, 2.1414], [-0.2815, 0.7482, 1.5920]])
    print("4 reference known answer (test_iou.py: 0.9944; fp64 half-space oracle 0.99411):",
          float(api.box3d_iou(c1.to(d), c2.to(d)[None])[0]))
    g = torch.Generator().manual_seed(0)
    n = 300
    cub = torch.zeros(n + 1, 15)
    cub[:, :3] = torch.rand(n + 1, 3, generator=g) + torch.tensor([0.0, 0.0, 3.0])
    cub[:, 3:6] = torch.rand(n + 1, 3, generator=g) * 1.5 + 0.2
    cub[:, 6:] = O.random_rotations(1, n + 1, g)[0].reshape(n + 1, 9)
    cub[1] = cub[0]
    corners = api.cube_corners(cub.to(d))
    iou = api.box3d_iou(corners[0], corners[1:]).cpu().numpy()
    ref = O.box3d_iou_np(corners[0].cpu().numpy(), corners[1:].cpu().numpy())
    print("4 random boxes: max abs err %.2e, identical -> %.7f" % (np.abs(iou - ref).max(), iou[0]))


if __name__ == "__main__":
    part1(); part4(); part3(); part2()
