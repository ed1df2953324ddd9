"""ProposalNetwork/proposals/proposals.py: the production sampler `propose` (:338-424) on the GPU.

Random numbers are drawn with torch in the reference's program order (R1..R7 of SURVEY.md 3.5: three truncated
normal tensors with their data-dependent redraws, three plain normal tensors, then the rotation draw from the CPU
generator), so a seeded call consumes torch's generators exactly like the reference and produces the same
proposal set; the arithmetic (diagonal depth gather, pinhole trig, median / std over the n seeds, scaling, yaw
table lookup) runs in one libtdod launch (csrc/fused.cu, propose_only mode)."""
import numpy as np
import torch

from tdod import api
from tdod import lib as L
from ProposalNetwork.utils import utils
from ProposalNetwork.utils.spaces import Cubes
from ProposalNetwork.utils.utils import gt_in_norm_range

# 0.0x meters is the minimum edge length (proposals.py:10)
MIN_PROP_S = 0.05


def rescale_interval(x, min, max):
    """proposals.py:12-14"""
    return (min - max) * x + max


def lin_fun(x, coef):
    """proposals.py:16-18"""
    return coef[0] * x + coef[1]


def _boxes_tensor(reference_box):
    t = reference_box.tensor
    if not t.is_cuda:
        raise ValueError("reference_box must live on a CUDA device: the B200 path has no CPU implementation")
    return t.to(torch.float32).contiguous()


RATIOS = [0.33, 0.66, 1, 1.33, 1.67, 2, 3]   # proposals.py:232


def _launch(boxes, depth_image, priors, K, n, noise, rot, centre_mode, dims_mode, im_shape=None, aux=None):
    """one libtdod launch (csrc/fused.cu, propose_only): noise [N,6,n] (w,h,l,x,y,z), rot [N,n,9] -> PssOutputs"""
    dev = boxes.device
    N = boxes.shape[0]
    # the callers pass None for what a variant does not use (roi_heads.py:286-297)
    if depth_image is None:
        if centre_mode == L.CENTRE_DEPTH:
            raise ValueError("this sampler needs a depth image")
        depth_image = torch.zeros((1, 1), dtype=torch.float32, device=dev)
    if K is None:
        if centre_mode == L.CENTRE_DEPTH:
            raise ValueError("this sampler needs K")
        K = torch.eye(3, dtype=torch.float32, device=dev)
    depth = depth_image.to(dev, torch.float32).contiguous()
    if depth.dim() != 2:
        raise ValueError("depth_image must be [H,W]")
    if priors is None:
        mean = std = torch.ones((N, 3), dtype=torch.float32, device=dev)
    else:
        mean = priors[0].to(dev, torch.float32).contiguous()
        std = priors[1].to(dev, torch.float32).contiguous()
    Kd = K.to(dev, torch.float32).reshape(1, 3, 3).contiguous()
    return api.propose_score_select(depth[None], Kd, boxes, mean, std, torch.zeros(N, dtype=torch.int32, device=dev),
                                    None, None, n, noise=noise.contiguous(), rot=rot, propose_only=True,
                                    centre_mode=centre_mode, dims_mode=dims_mode, im_shape=im_shape, aux=aux)


def _finish(o, gt_cubes, ones_ranges=True):
    cubes = Cubes(o.all_cubes)
    if gt_cubes is None:
        return cubes, None, None
    t = o.all_cubes
    stats = statistics(gt_cubes, t[..., 0], t[..., 1], t[..., 2], t[..., 3], t[..., 4], t[..., 5])
    return cubes, stats, torch.ones(cubes.num_instances, 9)


def _rot(n, N, dev):
    return utils.randn_orthobasis_torch(n, N).reshape(N, n, 9).to(dev).contiguous()


def _n_ok(number_of_proposals):
    n = int(number_of_proposals)
    if n < 2:
        raise ValueError("number_of_proposals must be >= 2")
    return n


def propose_random(reference_box, depth_image, priors, im_shape, K, number_of_proposals=1, gt_cubes=None,
                   ground_normal: torch.Tensor = None):
    """proposals.py:20-45: centres and dimensions uniform, random rotations.  Draw order: x, y, z, w, h, l, R."""
    boxes = _boxes_tensor(reference_box)
    dev, N, n = boxes.device, boxes.shape[0], _n_ok(number_of_proposals)
    u = [torch.rand(N, n, device=dev) for _ in range(6)]            # x y z w h l
    noise = torch.stack([u[3], u[4], u[5], u[0], u[1], u[2]], dim=1)
    o = _launch(boxes, depth_image, priors, K, n, noise, _rot(n, N, dev), L.CENTRE_UNIFORM, L.DIMS_UNIFORM)
    return _finish(o, gt_cubes)


def propose_xy_patch(reference_box, depth_image, priors, im_shape, K, number_of_proposals=1, gt_cubes=None,
                     ground_normal: torch.Tensor = None):
    """proposals.py:47-91: x, y on a ramp over the inner half of the box (normalised image space), z uniform,
    dimensions uniform.  Draw order: z, w, h, l, R."""
    boxes = _boxes_tensor(reference_box)
    dev, N, n = boxes.device, boxes.shape[0], _n_ok(number_of_proposals)
    zero = torch.zeros(N, n, device=dev)
    uz = torch.rand(N, n, device=dev)
    uw, uh, ul = (torch.rand(N, n, device=dev) for _ in range(3))
    noise = torch.stack([uw, uh, ul, zero, zero, uz], dim=1)
    o = _launch(boxes, depth_image, priors, K, n, noise, _rot(n, N, dev), L.CENTRE_PATCH, L.DIMS_UNIFORM, im_shape)
    return _finish(o, gt_cubes)


def propose_z(reference_box, depth_image, priors, im_shape, K, number_of_proposals=1, gt_cubes=None,
              ground_normal: torch.Tensor = None):
    """proposals.py:93-135: x, y as propose_xy_patch; z = torch.linspace between the 10 % and 90 % quantiles of the
    depth inside the box (:109-115, torch.quantile / torch.linspace per object -- host-side torch calls on the device,
    as in the reference); dimensions uniform.  Draw order: w, h, l, R."""
    boxes = _boxes_tensor(reference_box)
    dev, N, n = boxes.device, boxes.shape[0], _n_ok(number_of_proposals)
    depth = depth_image.to(dev, torch.float32)
    z = torch.zeros(N, n, device=dev)
    qs = torch.tensor([0.1, 0.9], device=dev)
    bi = boxes.cpu()
    for i in range(N):
        patch = depth[int(bi[i, 1]):int(bi[i, 3]), int(bi[i, 0]):int(bi[i, 2])]
        q = torch.quantile(patch, qs, dim=None)
        z[i] = torch.linspace(q[0], q[1], n)
    zero = torch.zeros(N, n, device=dev)
    uw, uh, ul = (torch.rand(N, n, device=dev) for _ in range(3))
    noise = torch.stack([uw, uh, ul, zero, zero, z], dim=1)
    o = _launch(boxes, depth, priors, K, n, noise, _rot(n, N, dev), L.CENTRE_PATCH_Z, L.DIMS_UNIFORM, im_shape)
    return _finish(o, gt_cubes)


def propose_random_dim(reference_box, depth_image, priors, im_shape, K, number_of_proposals=1, gt_cubes=None,
                       ground_normal: torch.Tensor = None):
    """proposals.py:137-197: centres as `propose`, dimensions uniform.  Draw order: w, h, l, x, y, z, R."""
    boxes = _boxes_tensor(reference_box)
    dev, N, n = boxes.device, boxes.shape[0], _n_ok(number_of_proposals)
    uw, uh, ul = (torch.rand(N, n, device=dev) for _ in range(3))
    zx, zy, zz = (torch.randn((N, n), device=dev) for _ in range(3))
    noise = torch.stack([uw, uh, ul, zx, zy, zz], dim=1)
    o = _launch(boxes, depth_image, priors, K, n, noise, _rot(n, N, dev), L.CENTRE_DEPTH, L.DIMS_UNIFORM)
    return _finish(o, gt_cubes)


def propose_aspect_ratio(reference_box, depth_image, priors, im_shape, K, number_of_proposals=1, gt_cubes=None,
                         ground_normal: torch.Tensor = None):
    """proposals.py:199-269: w uniform, h and l = w times one of seven fixed ratios chosen per object with
    torch.randperm(7)[0] (twice per object, CPU generator, :238); centres as `propose`.
    Draw order: w, (ratio_h, ratio_l) per object, x, y, z, R."""
    boxes = _boxes_tensor(reference_box)
    dev, N, n = boxes.device, boxes.shape[0], _n_ok(number_of_proposals)
    uw = torch.rand(N, n, device=dev)
    aux = torch.empty((N, 2), dtype=torch.float32)
    for i in range(N):
        r1, r2 = torch.randperm(len(RATIOS))[0], torch.randperm(len(RATIOS))[0]
        aux[i, 0], aux[i, 1] = RATIOS[r1], RATIOS[r2]
    zx, zy, zz = (torch.randn((N, n), device=dev) for _ in range(3))
    zero = torch.zeros(N, n, device=dev)
    noise = torch.stack([uw, zero, zero, zx, zy, zz], dim=1)
    o = _launch(boxes, depth_image, priors, K, n, noise, _rot(n, N, dev), L.CENTRE_DEPTH, L.DIMS_ASPECT,
                aux=aux.to(dev))
    return _finish(o, gt_cubes)


def propose_random_rotation(reference_box, depth_image, priors, im_shape, K, number_of_proposals=1, gt_cubes=None,
                            ground_normal: torch.Tensor = None):
    """proposals.py:272-336: `propose` with uniformly random rotations whatever the ground normal."""
    return propose(reference_box, depth_image, priors, im_shape, K, number_of_proposals, gt_cubes, None)


def propose(reference_box, depth_image, priors, im_shape, K, number_of_proposals=1, gt_cubes=None,
            ground_normal: torch.Tensor = None):
    """proposals.py:338-424.  reference_box: Boxes [N] (xyxy px); depth_image [H,W] fp32 metres;
    priors = (mean [N,3], std [N,3]) in (w,h,l) order; im_shape = (width, height); K [3,3].
    Returns (Cubes [N,n,15], stats | None, ranges | None)."""
    boxes = _boxes_tensor(reference_box)
    dev = boxes.device
    N, n = boxes.shape[0], int(number_of_proposals)
    if n < 2:
        raise ValueError("number_of_proposals must be >= 2 (the reference's linspace spacing divides by n-1)")
    if depth_image.dim() != 2:
        raise ValueError("depth_image must be [H,W]")
    depth = depth_image.to(dev, torch.float32).contiguous()
    mean = priors[0].to(dev, torch.float32).contiguous()
    std = priors[1].to(dev, torch.float32).contiguous()
    Kd = K.to(dev, torch.float32).reshape(1, 3, 3).contiguous()

    # R1..R3: truncated dimension draws (accepted standard normals); R4..R6: centre draws
    _, zw = utils._accepted_normals(mean[:, 0], std[:, 0], n, MIN_PROP_S, mean[:, 0] + 2 * std[:, 0])
    _, zh = utils._accepted_normals(mean[:, 1], std[:, 1] * 1.1, n, MIN_PROP_S, mean[:, 1] + 2.2 * std[:, 1])
    _, zl = utils._accepted_normals(mean[:, 2], std[:, 2], n, MIN_PROP_S, mean[:, 2] + 2 * std[:, 2])
    zx = torch.randn((N, n), device=dev)
    zy = torch.randn((N, n), device=dev)
    zz = torch.randn((N, n), device=dev)
    noise = torch.stack([zw, zh, zl, zx, zy, zz], dim=1).contiguous()          # [N,6,n]

    # R7: pose
    yaw_tab = yaw_idx = rot = None
    if ground_normal is None:
        rot = utils.randn_orthobasis_torch(n, N).reshape(N, n, 9).to(dev).contiguous()
    else:
        gn = ground_normal.to(dev)
        if gn.dtype not in (torch.float32, torch.float64):
            gn = gn.to(torch.float32)
        yaw_tab = api.yaw_table(gn.reshape(1, 3).contiguous())
        yaw_idx = torch.randint(utils.N_YAW, (N, n)).to(torch.int16).to(dev)

    o = api.propose_score_select(depth[None], Kd, boxes, mean, std, torch.zeros(N, dtype=torch.int32, device=dev),
                                 None, None, n, yaw_tab=yaw_tab, noise=noise, yaw_idx=yaw_idx, rot=rot,
                                 propose_only=True)
    cubes = Cubes(o.all_cubes)
    if gt_cubes is None:
        return cubes, None, None

    t = o.all_cubes
    stats = statistics(gt_cubes, t[..., 0], t[..., 1], t[..., 2], t[..., 3], t[..., 4], t[..., 5])
    m = gt_cubes.num_instances
    pi = torch.full((m,), float(torch.pi), device=dev)
    # o.stats[:, 1::2] = std(x)*1.2, std(y)*0.8, std(z)*1.2 as used by the samplers (proposals.py:420)
    ranges = torch.stack([o.stats[:, 1], o.stats[:, 3], o.stats[:, 5], std[:, 0], std[:, 1] * 1.1, std[:, 2], pi, pi,
                          pi], dim=1).cpu().numpy()
    return cubes, stats, ranges


def statistics(gt_cubes, x, y, z, w, h, l):
    """proposals.py:427-444 (MABO bookkeeping; host loop over the few GT objects, min/max on the device)"""
    from cubercnn import util  # mat2euler lives in the caller's package (proposals.py:7)
    m = gt_cubes.num_instances
    stats = torch.zeros((m, 9))
    lo = [v.min(dim=1).values.cpu() for v in (x, y, z, w, h, l)]
    hi = [v.max(dim=1).values.cpu() for v in (x, y, z, w, h, l)]
    for i in range(m):
        gt = gt_cubes[i].tensor[0, 0].cpu()
        s = [gt_in_norm_range([lo[k][i], hi[k][i]], gt[k]) for k in range(6)]
        ang = util.mat2euler(gt[-9:].reshape((3, 3)))
        s.append(gt_in_norm_range(torch.tensor([0, np.pi]), torch.tensor(ang[0])))
        s.append(gt_in_norm_range(torch.tensor([0, np.pi / 2]), torch.tensor(ang[1])))
        s.append(gt_in_norm_range(torch.tensor([0, np.pi]), torch.tensor(ang[2])))
        stats[i] = torch.tensor(s)
    return stats
