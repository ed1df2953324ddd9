"""Round-2 GPU probe (not a test): numbers that decide which parity gates can be asserted.

  A. the oracle's torch code run on CUDA TENSORS (the reference as it runs in production, proposals.py:354-396 on
     reference_box.device) against the fused kernel in HOST-RNG mode: seeds, medians, stds, centres, scores
  B. PHILOX mode replayed through HOST mode from the kernel's own out_noise / out_yaw_idx; distribution of the draws
  C. tdod_box3d_iou against the fp64 half-space oracle and the reference's known answer
  D. frames without a ground map: plane on all strided points (roi_heads.py:358-376)

    python tests/probes/r2_probe1.py > gpurun_out/r2_probe1.txt
"""
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


def ulps(a, b):
    a = a.contiguous().view(torch.int32).long(); b = b.contiguous().view(torch.int32).long()
    return (a - b).abs()


def part_a(frame_ids, cfg, n):
    c = synth.CONFIGS[cfg]
    tot = dict(objs=0, seeds_ne=0, med_ne=0, std_ne=0, std_maxulp=0, ctr_ne=0, ctr_ne_objs=0, ctr_maxulp=0, ctr_own_stats_ne=0,
               whlR_ne=0, sc_miss_gpu={}, sc_miss_cpu={}, sc_n=0, win_ne_gpu=0, win_ne_cpu=0, win_checked=0)
    for fid in frame_ids:
        f = synth.make_frame(fid, c["H"], c["W"], c["n_obj"])
        N = f.n_obj()
        if N == 0:
            continue
        H, W = c["H"], c["W"]
        pg, _ = O.backproject_np(f.depth.numpy(), f.K.numpy(), f.ground.numpy())
        ids = O.draw_sample_ids(len(pg), 1000, random.Random(fid))
        eq, _, _ = O.ransac_np(pg, ids, 0.05)
        nv = torch.from_numpy(O.rectify_normal(np.array(eq[:3])))
        o = O.propose(f.boxes, f.depth, f.prior_mean, f.prior_std, f.K, n, nv, torch.Generator().manual_seed(fid))
        noise = o["noise"]                                        # [6,N,n]
        # oracle on CUDA tensors
        g = lambda t: t.to(d)
        cub_gpu = O.cubes_from_noise(g(f.boxes), g(f.depth), g(f.prior_mean), g(f.prior_std), g(f.K), n, g(noise), nv,
                                     o["yaw_idx"])
        x0, y0, z0 = O.centre_seeds(g(f.boxes), g(f.depth), g(f.K), n)
        zt = z0 + cub_gpu[:, :, 5] / 2
        seeds_ref = torch.stack([x0, y0, zt], 1)                  # [N,3,n]
        med_ref = torch.stack([torch.median(s, dim=1).values for s in (x0, y0, zt)], 1)
        std_ref = torch.stack([torch.std(s, dim=1) for s in (x0, y0, zt)], 1)
        # kernel
        rects = torch.zeros((N, 4, 2), device=d)
        m_rect, has = api.mask_rect(g(f.masks))
        yaw_tab = api.yaw_table(nv.to(d)[None])
        k = api.propose_score_select(g(f.depth)[None], g(f.K)[None], g(f.boxes), g(f.prior_mean), g(f.prior_std),
                                     torch.zeros(N, dtype=torch.int32, device=d), m_rect, has, n, yaw_tab=yaw_tab,
                                     noise=g(noise.permute(1, 0, 2).contiguous()), yaw_idx=g(o["yaw_idx"].to(torch.int16)),
                                     return_all=True, want_seeds=True)
        torch.cuda.synchronize()
        tot["objs"] += N
        tot["seeds_ne"] += int((ulps(k.seeds, seeds_ref) != 0).sum())
        coef_a = torch.tensor([1.15, 1.1, 0.85], device=d); coef_b = torch.tensor([0.0, 0.0, 0.35], device=d)
        coef_s = torch.tensor([1.2, 0.8, 1.2], device=d)
        mean_ref = coef_a * med_ref + coef_b
        sd_ref = std_ref * coef_s
        st = k.stats.reshape(N, 3, 2)
        tot["med_ne"] += int((ulps(st[:, :, 0], mean_ref) != 0).sum())
        us = ulps(st[:, :, 1], sd_ref)
        tot["std_ne"] += int((us != 0).sum()); tot["std_maxulp"] = max(tot["std_maxulp"], int(us.max()))
        uc = ulps(k.all_cubes[:, :, :3], cub_gpu[:, :, :3])
        tot["ctr_ne"] += int((uc != 0).sum()); tot["ctr_maxulp"] = max(tot["ctr_maxulp"], int(uc.max()))
        tot["ctr_ne_objs"] += int(((uc != 0).sum((1, 2)) > 0).sum())
        tot["whlR_ne"] += int((ulps(k.all_cubes[:, :, 3:], cub_gpu[:, :, 3:]) != 0).sum())
        # centres from the kernel's own (mean, sd): only the z*sd+mean chain
        own = torch.stack([g(noise[3 + a]) * st[:, a, 1:2] + st[:, a, 0:1] for a in range(3)], 2)
        tot["ctr_own_stats_ne"] += int((ulps(k.all_cubes[:, :, :3], own) != 0).sum())
        # scores: oracle on CUDA tensors for iou / dims; corners via cKDTree needs the host (fp64, exact NN)
        for flavour, cubes_ref in (("gpu", cub_gpu), ("cpu", o["cubes"])):
            cr = cubes_ref.to(d) if flavour == "gpu" else cubes_ref
            Kt = g(f.K) if flavour == "gpu" else f.K
            c2d = O.project_corners(O.cube_corners(cr), Kt, W, H)
            pb = O.boxes_from_corners(c2d)
            miss = tot["sc_miss_" + flavour]
            for i in range(N):
                bx = (g(f.boxes[i]) if flavour == "gpu" else f.boxes[i])
                pm = (g(f.prior_mean[i]) if flavour == "gpu" else f.prior_mean[i])
                ps = (g(f.prior_std[i]) if flavour == "gpu" else f.prior_std[i])
                s_iou = O.iou_score(bx, pb[i]).cpu()
                s_dim = O.dims_score(pm, ps, cr[i, :, 3:6], bx, pb[i]).cpu()
                rect = m_rect[i].cpu().numpy() if int(has[i]) == 1 else None
                s_cor = O.corners_score(f.masks[i], c2d[i].cpu(), rect=rect if rect is not None else None)
                comb, best = O.select(s_iou, s_dim, s_cor)
                got = k.all_scores[i].cpu().numpy()
                for ch, ref in enumerate((s_iou.numpy(), s_dim.numpy(), s_cor.numpy(), comb)):
                    bad = np.abs(got[ch] - ref) > 1e-7 + 1e-5 * np.abs(ref)
                    bad &= ~(np.isnan(ref) & np.isnan(got[ch]))
                    miss[ch] = miss.get(ch, 0) + int(bad.sum())
                if flavour == "gpu":
                    tot["sc_n"] += n
                srt = np.sort(comb[~np.isnan(comb)])
                if not np.isnan(comb).any() and len(srt) > 1 and srt[-1] - srt[-2] > 1e-5 * abs(srt[-1]):
                    if flavour == "gpu":
                        tot["win_checked"] += 1
                    tot["win_ne_" + flavour] += int(int(k.best_idx[i]) != best)
    return tot


def part_b():
    frames = synth.make_frames("C2", range(16))
    b = pipeline.FrameBatch.from_frames(frames, d)
    s = pipeline.select(b, 1000, seed=7, want_noise=True, return_all=True)
    torch.cuda.synchronize()
    o = s.all
    yaw_tab = api.yaw_table(s.normal)
    k = api.propose_score_select(b.depth, b.K, b.boxes, b.prior_mean, b.prior_std, b.obj_frame, s.rects, s.has_rect, 1000,
                                 yaw_tab=yaw_tab, noise=o.noise, yaw_idx=o.yaw_idx, return_all=True)
    torch.cuda.synchronize()
    eq = lambda a, c: bool(((a == c) | (a.isnan() & c.isnan())).all())
    print("B replay PHILOX -> HOST: cubes", eq(k.all_cubes, o.all_cubes), "boxes", eq(k.all_boxes, o.all_boxes), "scores",
          eq(k.all_scores, o.all_scores), "best_idx", bool(torch.equal(k.best_idx, o.best_idx)))
    # distribution of the draws behind the centres (x,y,z: plain normals) and of the yaw bins
    z = o.noise[:, 3:6].double().flatten()
    nz = z.numel()
    print("B normals n=%d mean %.5f var %.5f skew %.5f kurt %.5f  tail>3s %.3e (2.700e-03)  tail>4s %.3e (6.334e-05)" % (
        nz, z.mean(), z.var(), (z ** 3).mean(), (z ** 4).mean(), float((z.abs() > 3).double().mean()),
        float((z.abs() > 4).double().mean())))
    zs = torch.sort(z[torch.randperm(nz, device=z.device)[:200000]]).values
    cdf = 0.5 * (1 + torch.erf(zs / 2 ** 0.5))
    emp = (torch.arange(1, zs.numel() + 1, device=z.device).double()) / zs.numel()
    ks = float((cdf - emp).abs().max())
    print("B KS statistic over 200k draws %.5f (1%% critical value %.5f)" % (ks, 1.63 / zs.numel() ** 0.5))
    yi = o.yaw_idx.flatten().long()
    cnt = torch.bincount(yi, minlength=36).double()
    chi2 = float(((cnt - cnt.mean()) ** 2 / cnt.mean()).sum())
    print("B yaw bins chi2 %.1f (35 dof, 1%% critical 57.3) min %d max %d" % (chi2, int(cnt.min()), int(cnt.max())))
    w = o.noise[:, 0:3]
    print("B truncated dims: all within [0.05, mean+k*std]:",
          bool((o.all_cubes[:, :, 3:6] >= 0.05).all()))


def part_c():
    g = torch.Generator().manual_seed(0)
    n = 400
    cub = torch.zeros(n + 1, 15)
    cub[:, :3] = torch.rand(n + 1, 3, generator=g) * torch.tensor([1.0, 1.0, 1.0]) + torch.tensor([0.0, 0.0, 3.0])
    cub[:, 3:6] = torch.rand(n + 1, 3, generator=g) * 1.5 + 0.2
    R = O.random_rotations(1, n + 1, g)[0]
    cub[:, 6:] = R.reshape(n + 1, 9)
    cub[1] = cub[0]                                           # identical box -> 1
    cub[2] = cub[0]; cub[2, 0] += 50.0                        # disjoint -> 0
    cub[3] = cub[0]; cub[3, 3:6] *= 0.5                       # nested -> 1/8
    cub[4, 6:] = torch.eye(3).reshape(9); cub[5] = cub[4]; cub[5, 0] += cub[4, 5]   # touching faces (axis X <-> l) -> 0
    corners = api.cube_corners(cub.to(d))
    iou = api.box3d_iou(corners[0], corners[1:]).cpu().numpy()
    ref = O.box3d_iou_np(corners[0].cpu().numpy(), corners[1:].cpu().numpy())
    err = np.abs(iou - ref)
    print("C box3d_iou vs fp64 half-space oracle: n=%d max abs err %.2e mean %.2e; identical %.7f disjoint %.7f nested %.7f" % (
        n, err.max(), err.mean(), iou[0], iou[1], iou[2]), "nonzero", int((ref > 0).sum()))
    c45 = api.cube_corners(cub[4:6].to(d))
    print("C touching faces:", float(api.box3d_iou(c45[0], c45[1:])[0]))
    c1 = torch.tensor([[0.2411, -0.1752, 1.2247], [0.1951, -0.4194, 1.7741], [0.2036, 0.4826, 2.1757], [0.2495, 0.7267, 1.6263],
                       [-0.2920, -0.1549, 1.1903], [-0.3380, -0.3991, 1.7396], [-0.3295, 0.5029, 2.1412], [-0.2835, 0.7471, 1.5919]])
    c2 = torch.tensor([[0.2390, -0.1764, 1.2246], [0.1930, -0.4205, 1.7740], [0.2055, 0.4813, 2.1759], [0.2515, 0.7254, 1.6265],
                       [-0.2940, -0.1536, 1.1901], [-0.3400, -0.3978, 1.7395], [-0.3274, 0.5040, 2.1414], [-0.2815, 0.7482, 1.5920]])
    print("C reference known answer (test_iou.py: 0.9944):", float(api.box3d_iou(c1.to(d), c2.to(d)[None])[0]))


def part_d():
    c = synth.CONFIGS["C1"]
    bad = 0
    for fid in range(6):
        f = synth.make_frame(fid, c["H"], c["W"], c["n_obj"])
        f.ground = None
        pts, _ = O.backproject_np(f.depth.numpy(), f.K.numpy(), None)
        ids = O.draw_sample_ids(len(pts), 1000, random.Random(fid))
        eq, inl, _ = O.ransac_np(pts, ids, 0.05)
        nv = O.rectify_normal(np.array(eq[:3]))
        b = pipeline.FrameBatch.from_frames([f], d)
        s = pipeline.select(b, 256, seed=3, sample_ids=torch.from_numpy(ids.astype(np.int32))[None].contiguous().to(d))
        torch.cuda.synchronize()
        ok = int(s.inlier_count[0]) == len(inl) and np.allclose(s.normal[0].cpu().numpy(), nv, rtol=0, atol=1e-12)
        bad += not ok
    print("D no-ground frames: %d / 6 planes differ from the oracle" % bad)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "ABCD"
    if "C" in which:
        part_c()
    if "D" in which:
        part_d()
    if "B" in which:
        part_b()
    if "A" in which:
        print("A C1 frames 0..5:", part_a(range(6), "C1", 1000))
        print("A C2 frames 0..5:", part_a(range(6), "C2", 1000))
