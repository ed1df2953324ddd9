"""Batched proposal-and-scoring over many frames: the fast entry SURVEY.md 8(b) calls an *addition* to the
per-object drop-in functions.  One call covers what `ROIHeads_Boxer._forward_cube`
(cubercnn/modeling/roi_heads/roi_heads.py:304-505) does frame by frame and object by object:

    depth -> strided point cloud -> RANSAC ground plane -> y-up normal      (roi_heads.py:345-428)   K-A
    normal -> 36-entry yaw table                                           (proposals.py:404-407)
    object mask -> min-area rectangle of its largest contour               (scorefunction.py:59-68) K-B
    propose + project + IoU x dims x corners + argmax, fused               (roi_heads.py:493-505)   K-C

Everything runs on the GPU through libtdod; there is no CPU path.  `FrameBatch` holds device tensors,
`HostBatch` pinned host tensors; `Pipeline.run_host` streams a HostBatch through the device in chunks of
frames with the copies of chunk k+1 overlapping the kernels of chunk k.
"""
import os

import torch

from . import api
from . import lib as L

f32, i32, u8 = torch.float32, torch.int32, torch.uint8


class FrameBatch:
    """F frames and their N objects (CSR over frames), all on one device.

    depth [F,H,W] f32 · ground [F,H,W] u8/bool or None · K [F,3,3] f32 · obj_off [F+1] (python list / CPU tensor)
    boxes [N,4] f32 xyxy · prior_mean/prior_std [N,3] f32 (w,h,l) · obj_frame [N] i32 · masks [N,H,W] u8/bool"""
    __slots__ = ("depth", "ground", "K", "boxes", "prior_mean", "prior_std", "obj_frame", "masks", "frame_ids",
                 "has_ground", "masks_packed")

    def __init__(self, depth, ground, K, boxes, prior_mean, prior_std, obj_frame, masks, frame_ids=None,
                 has_ground=None, masks_packed=None):
        self.depth, self.ground, self.K = depth, ground, K
        self.boxes, self.prior_mean, self.prior_std = boxes, prior_mean, prior_std
        self.obj_frame, self.masks, self.frame_ids = obj_frame, masks, frame_ids
        # per-frame python bools, or None = every frame has a ground map (when `ground` is not None).  A frame without one
        # carries an all-ones map: `ground > 0` then selects every strided point, which is what the reference fits the plane
        # on when ground_maps is None (roi_heads.py:358-371, dataset_mapper.py:146)
        self.has_ground = has_ground
        # optional alternative to `masks` (then None): the masks at 1 bit per pixel, int32 [N,H,ceil(W/32)], see `pack_masks`
        # -- an eighth of the bytes to keep in host memory and to move over PCIe
        self.masks_packed = masks_packed

    @property
    def n_frames(self):
        return int(self.depth.shape[0])

    @property
    def n_obj(self):
        return int(self.boxes.shape[0])

    @property
    def device(self):
        return self.depth.device

    def to(self, device, non_blocking=False):
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)
        return FrameBatch(mv(self.depth), mv(self.ground), mv(self.K), mv(self.boxes), mv(self.prior_mean),
                          mv(self.prior_std), mv(self.obj_frame), mv(self.masks), self.frame_ids, self.has_ground,
                          mv(self.masks_packed))

    def pin(self):
        pm = lambda t: None if t is None else t.pin_memory()
        return FrameBatch(pm(self.depth), pm(self.ground), pm(self.K), pm(self.boxes), pm(self.prior_mean),
                          pm(self.prior_std), pm(self.obj_frame), pm(self.masks), self.frame_ids, self.has_ground,
                          pm(self.masks_packed))

    def input_bytes(self):
        ts = (self.depth, self.ground, self.K, self.boxes, self.prior_mean, self.prior_std, self.obj_frame, self.masks,
              self.masks_packed)
        return sum(t.numel() * t.element_size() for t in ts if t is not None)

    @staticmethod
    def from_frames(frames, device=None):
        """stack `tdod.synth.Frame`s (or anything with the same fields) into one batch"""
        dev = frames[0].depth.device if device is None else torch.device(device)
        counts = [f.n_obj() for f in frames]
        obj_frame = torch.repeat_interleave(torch.arange(len(frames), dtype=i32), torch.tensor(counts)).to(dev)
        cat = lambda name: torch.cat([getattr(f, name) for f in frames if f.n_obj() > 0]).to(dev)
        masks = None
        if frames[0].masks is not None:
            masks = torch.cat([f.masks for f in frames if f.n_obj() > 0]).to(dev)
        has = [f.ground is not None for f in frames]
        ground, has_ground = None, None
        if any(has):
            proto = next(f.ground for f in frames if f.ground is not None)
            ground = torch.stack([f.ground if f.ground is not None else torch.ones_like(proto) for f in frames]).to(dev)
            has_ground = None if all(has) else has
        return FrameBatch(torch.stack([f.depth for f in frames]).to(dev), ground,
                          torch.stack([f.K for f in frames]).to(dev), cat("boxes"), cat("prior_mean"),
                          cat("prior_std"), obj_frame, masks, [f.frame_id for f in frames], has_ground)


def pack_masks(masks):
    """masks [N,H,W] bool / uint8 (any device) -> int32 [N,H,ceil(W/32)]: bit (c & 31) of word (c >> 5) of a row = pixel c
    (the layout tdod_mask_rect_bitmap reads).  Host-side convenience for callers that want to keep or ship masks packed."""
    m = (masks != 0)
    N, H, W = m.shape
    Wp = (W + 31) // 32
    pad = Wp * 32 - W
    if pad:
        m = torch.nn.functional.pad(m, (0, pad))
    w = torch.tensor([1 << k for k in range(31)] + [-(1 << 31)], dtype=torch.int64, device=m.device)
    return (m.reshape(N, H, Wp, 32).to(torch.int64) * w).sum(-1).to(torch.int32)


def with_packed_masks(batch):
    """the same batch with its masks at 1 bit per pixel only"""
    return FrameBatch(batch.depth, batch.ground, batch.K, batch.boxes, batch.prior_mean, batch.prior_std, batch.obj_frame,
                      None, batch.frame_ids, batch.has_ground, pack_masks(batch.masks))


class Selection:
    """winners per object (+ per-frame plane outputs)"""
    __slots__ = ("best_idx", "best_score", "best_cube", "best_box", "normal", "plane_eq", "inlier_count", "rects",
                 "has_rect", "all")


def ground_normals(depth, ground, K, seed=0, sample_ids=None, iters=1000, thresh=0.05, stride=5,
                   flavour=L.BACKPROJECT_NUMPY, want_other=False, frame_id_base=0):
    """K-A for a stack of frames: (normal [F,3], eq [F,4], count [F], plane_pts, other_pts, n_pts [F,2]).
    flavour NUMPY = the fp64 path ROIHeads_Boxer uses (roi_heads.py:345-428), TORCH = the fp32 path of
    ROIHeads3DScore.normal_vector_from_maps (:1084-1136)."""
    pts, other, counts = api.backproject(depth, ground, K, stride, flavour, want_other=want_other)
    r = api.ransac_plane(pts, counts[:, 0].contiguous(), sample_ids, seed=seed, iters=iters, thresh=thresh,
                         frame_id_base=frame_id_base)
    return r["normal"], r["eq"], r["count"], pts, other, counts


def draw_reference_noise(batch, obj_off, n_prop, generator=None, cpu_generator=None, with_normal=True):
    """The random numbers of `propose` (ProposalNetwork/proposals/proposals.py:377-407) drawn WITH TORCH'S GENERATORS IN THE
    REFERENCE'S PROGRAM ORDER, frame by frame as the reference calls it: per frame three truncated draws for w, h, l --
    each a whole-tensor redraw loop over the frame's [N_f, n] block until no entry is out of range
    (sample_normal_in_range, ProposalNetwork/utils/utils.py:42-60) -- then x, y, z, all from the device generator; the
    yaw bins from `torch.randint` on the CPU generator (the reference builds that index tensor on the host, :407).
    Returns (noise [N,6,n] accepted standard normals, yaw_idx [N,n] int16 | None): what HOST mode consumes, so that a
    seeded run reproduces the reference's proposal set (with std_mode TORCH_CUDA: bit for bit).
    The redraw loop synchronises with the host once per pass, as the reference's does."""
    dev = batch.device
    N = batch.n_obj
    noise = torch.empty((N, 6, n_prop), dtype=f32, device=dev)
    yaw = torch.empty((N, n_prop), dtype=torch.int16) if with_normal else None
    mu, sg = batch.prior_mean, batch.prior_std
    hs = torch.tensor([1.0, 1.1, 1.0], device=dev)
    hk = torch.tensor([2.0, 2.2, 2.0], device=dev)
    for f in range(batch.n_frames):
        o0, o1 = obj_off[f], obj_off[f + 1]
        if o1 == o0:
            continue
        nf = o1 - o0
        for k in range(3):
            mean, std = mu[o0:o1, k], sg[o0:o1, k] * hs[k]
            hi = (mu[o0:o1, k] + hk[k] * sg[o0:o1, k])[:, None]
            z = torch.randn((nf, n_prop), generator=generator, device=dev)
            v = z * std[:, None] + mean[:, None]
            tries = 0
            while True:
                bad = (v < 0.05) | (v > hi)
                if not bool(torch.any(bad)):
                    break
                z2 = torch.randn((nf, n_prop), generator=generator, device=dev)
                v = torch.where(bad, z2 * std[:, None] + mean[:, None], v)
                z = torch.where(bad, z2, z)
                tries += 1
                if tries == 10000:
                    break
            noise[o0:o1, k] = z
        for k in range(3, 6):
            noise[o0:o1, k] = torch.randn((nf, n_prop), generator=generator, device=dev)
        if with_normal:
            yaw[o0:o1] = torch.randint(36, (nf, n_prop), generator=cpu_generator).to(torch.int16)
    return noise, (yaw.to(dev, non_blocking=True) if with_normal else None)


_SIDE = {}


def _side_stream(dev):
    """one auxiliary stream per device: K-A (ALU-bound) runs there while K-B (HBM-bound) runs on the caller's stream"""
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
    if key not in _SIDE:
        _SIDE[key] = torch.cuda.Stream(dev)
    return _SIDE[key]


_SPLIT = {}


def _split_streams(dev, k):
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
    pool = _SPLIT.setdefault(key, [])
    while len(pool) < k:
        pool.append(torch.cuda.Stream(dev))
    return pool[:k]


def _select_split(batch, n_prop, seed, yaw_tab, side, main, obj_id_base, cluster_size, iou_boxes, splits):
    """K-B and K-C of `splits` object ranges on as many streams (optional; the latency-bound rectangle passes of one
    range can run beside the fused kernel and the pack of another -- on C2 the smaller launches cost more than that gains).  Results are written into slices of the
    same output tensors; Philox keys are global object ids, so the winners equal the unsplit run's."""
    dev = batch.device
    N = batch.n_obj
    o = api.PssOutputs()
    o.best_idx = torch.empty(N, dtype=i32, device=dev)
    o.best_score = torch.empty(N, dtype=f32, device=dev)
    o.best_cube = torch.empty((N, 15), dtype=f32, device=dev)
    o.best_box = torch.empty((N, 4), dtype=f32, device=dev)
    o.all_cubes = o.all_boxes = o.all_scores = o.stats = None
    rects = torch.empty((N, 4, 2), dtype=f32, device=dev)
    has = torch.empty(N, dtype=u8, device=dev)
    masks = batch.masks.view(u8) if batch.masks.dtype == torch.bool else batch.masks
    bounds = [N * g // splits for g in range(splits + 1)]
    streams = [main] + _split_streams(dev, splits - 1)
    for g, st in enumerate(streams):
        a, b = bounds[g], bounds[g + 1]
        if a == b:
            continue
        if st is not main:
            st.wait_stream(main)
        with torch.cuda.stream(st):
            api.mask_rect(masks[a:b], out=(rects[a:b], has[a:b]))
            if side is not None and side is not st:
                st.wait_stream(side)
            if yaw_tab is not None:
                yaw_tab.record_stream(st)
            og = api.PssOutputs()
            og.best_idx, og.best_score = o.best_idx[a:b], o.best_score[a:b]
            og.best_cube, og.best_box = o.best_cube[a:b], o.best_box[a:b]
            og.all_cubes = og.all_boxes = og.all_scores = og.stats = None
            api.propose_score_select(batch.depth, batch.K, batch.boxes[a:b], batch.prior_mean[a:b], batch.prior_std[a:b],
                                     batch.obj_frame[a:b], rects[a:b], has[a:b], n_prop, yaw_tab=yaw_tab, seed=seed,
                                     obj_id_base=obj_id_base + a, cluster_size=cluster_size, out=og,
                                     iou_boxes=None if iou_boxes is None else iou_boxes[a:b])
    for st in streams[1:]:
        main.wait_stream(st)
    return o, rects, has


def select(batch, n_prop, seed=0, noise=None, yaw_idx=None, rot=None, sample_ids=None, use_ground=True,
           ransac_iters=1000, ransac_thresh=0.05, flavour=L.BACKPROJECT_NUMPY, obj_id_base=0, frame_id_base=0,
           return_all=False, cluster_size=0, overlap=True, marks=None, iou_boxes=None, splits=None, want_noise=False,
           want_seeds=False, mode_b=False, std_mode=L.STD_EXACT):
    """Mode A (roi_heads.py:493-505) for every object of a device-resident batch.  Throughput RNG (in-kernel
    Philox keyed by global object id) unless `noise` (+ `yaw_idx` / `rot`) hands in host-drawn numbers.
    Launches only; does not synchronise.

    K-A (depth -> plane -> yaw table) and K-B (masks -> rectangles) are independent until K-C needs both: with
    `overlap` K-A is launched on an auxiliary stream so that its FP-bound hypothesis scoring shares the SMs with
    K-B's bandwidth-bound mask pass.  `marks`: optional dict that receives CUDA events
    (kb0/kb1 around K-B and kc0/kc1 around K-C on the caller's stream, ka0/ka1 around K-A on its stream).

    mode_b: also the three scores of the MABO branch (roi_heads.py:535-538) for every proposal -- segmentation and
    mod-segmentation IoU against the object's mask, points of the frame's non-ground cloud inside the cube's intervals --
    in the same fused call (`s.all.all_scores` channels 4, 5 and `s.all.all_pc`); implies return_all."""
    s = Selection()
    yaw_tab = None
    s.normal = s.plane_eq = s.inlier_count = None
    main = torch.cuda.current_stream(batch.device)

    def mark(name, stream):
        if marks is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream)
            marks[name] = ev

    # roi_heads.py:358-376: the plane is fitted whether or not the frame has a ground map (all strided points when it has
    # none) and its normal always reaches `propose`; use_ground=False is the ablation without a normal (random rotations)
    with_plane = bool(use_ground)
    side = _side_stream(batch.device) if (with_plane and overlap) else main
    if with_plane:
        if side is not main:
            side.wait_stream(main)
        with torch.cuda.stream(side):
            mark("ka0", side)
            s.normal, s.plane_eq, s.inlier_count, pts_plane, pts_other, pt_counts = ground_normals(
                batch.depth, batch.ground, batch.K, seed=seed, sample_ids=sample_ids, iters=ransac_iters,
                thresh=ransac_thresh, flavour=flavour, frame_id_base=frame_id_base, want_other=mode_b)
            yaw_tab = api.yaw_table(s.normal)
            if mode_b:
                # roi_heads.py:371-384: the cloud the point score sees is the non-ground points, or every point when the
                # frame has no ground map
                if batch.ground is None:
                    cloud, cloud_n = pts_plane, pt_counts[:, 0].contiguous()
                elif batch.has_ground is None:
                    cloud, cloud_n = pts_other, pt_counts[:, 1].contiguous()
                else:
                    hg = torch.tensor(batch.has_ground, device=batch.device)
                    cloud = torch.where(hg[:, None, None], pts_other, pts_plane)
                    cloud_n = torch.where(hg, pt_counts[:, 1], pt_counts[:, 0]).contiguous()
            mark("ka1", side)
    if splits is None:       # measured on C2 (B200): two ranges in flight 1.42 ms vs one range 1.34 ms -> off by default
        splits = 1
    if splits > 1:
        if noise is not None or return_all or mode_b:
            raise ValueError("splits > 1 needs the throughput RNG and select mode (no return_all)")
        if with_plane and side is not main:
            for t in (s.normal, s.plane_eq, s.inlier_count, yaw_tab):
                t.record_stream(main)
        o, s.rects, s.has_rect = _select_split(batch, n_prop, seed, yaw_tab, side if with_plane else None, main,
                                               obj_id_base, cluster_size, iou_boxes, splits)
        s.best_idx, s.best_score, s.best_cube, s.best_box = o.best_idx, o.best_score, o.best_cube, o.best_box
        s.all = None
        return s
    mark("kb0", main)
    if batch.masks is None and batch.masks_packed is not None:
        if mode_b:
            raise ValueError("mode_b samples the masks at 1 byte per pixel: pass `masks`, not `masks_packed`")
        s.rects, s.has_rect = api.mask_rect_bitmap(batch.masks_packed, int(batch.depth.shape[2]))
    else:
        s.rects, s.has_rect = api.mask_rect(batch.masks)
    mark("kb1", main)
    if with_plane and side is not main:
        main.wait_stream(side)
        for t in (s.normal, s.plane_eq, s.inlier_count, yaw_tab):
            t.record_stream(main)        # allocated on the auxiliary stream, consumed on the caller's
    if mode_b and not with_plane:
        raise ValueError("mode_b scores the frame's point cloud: use_ground must be on")
    if mode_b and side is not main:
        cloud.record_stream(main); cloud_n.record_stream(main)
    mark("kc0", main)
    o = api.propose_score_select(batch.depth, batch.K, batch.boxes, batch.prior_mean, batch.prior_std,
                                 batch.obj_frame, s.rects, s.has_rect, n_prop, yaw_tab=yaw_tab, noise=noise,
                                 yaw_idx=yaw_idx, rot=rot, seed=None if noise is not None else seed,
                                 obj_id_base=obj_id_base, return_all=return_all, cluster_size=cluster_size,
                                 iou_boxes=iou_boxes, want_noise=want_noise, want_seeds=want_seeds, std_mode=std_mode,
                                 masks=batch.masks if mode_b else None, points=cloud if mode_b else None,
                                 n_points=cloud_n if mode_b else None)
    mark("kc1", main)
    s.best_idx, s.best_score, s.best_cube, s.best_box = o.best_idx, o.best_score, o.best_cube, o.best_box
    s.all = o if (return_all or want_noise or want_seeds or mode_b) else None
    return s


# kernels launched by one `select` call: K-A 5 (back-projection, hypotheses, count, best, yaw table); K-B 5 (front, tail,
# list-driven pack, two exact passes -- the last three find empty work lists on ordinary masks); K-C 1
LAUNCHES_PER_SELECT = 11
LAUNCHES_PER_SELECT_NO_GROUND = 6


def launches_per_select(n_obj, with_plane=True, splits=None):
    """libtdod kernels one `select` call launches (masks that fit the fused K-B path)"""
    if splits is None:
        splits = 1
    return (5 if with_plane else 0) + 6 * splits


class GraphedSelect:
    """`select` on a resident FrameBatch captured ONCE into a CUDA graph (the 11 kernel nodes of a step, K-A forked onto
    the auxiliary stream inside the capture) and replayed: one driver call per step instead of eleven launches and their
    Python, so the GPU never waits for the host.  The graph reads the batch's buffers in place (refill them with copy_ to
    process new frames of the same shape) and writes the same output tensors at every replay.

    depth > 1 keeps `depth` independent copies of the graph (own workspaces and outputs) on `depth` streams and replays
    them round-robin: consecutive steps overlap, so the HBM-bound mask kernel of step i+1 runs beside the issue-bound
    proposal kernel of step i (software pipelining; 0.87 -> 0.82 ms per step on C2 at depth 2).  Each replay is ordered
    after what the caller's stream holds at that moment; `join()` orders the caller's stream after all replays.  With
    depth > 1 a step may still be reading the batch when `replay()` returns control to the caller's stream: `join()`
    before refilling the batch in place or reading an output.

        gs = GraphedSelect(batch, 1000, seed=7)
        out = gs.replay()            # depth 1: out is valid once the current stream reaches this point
    """

    def __init__(self, batch, n_prop, warmup=3, depth=1, **kw):
        dev = batch.device
        self.batch = batch
        self.depth = max(int(depth), 1)
        self.launches = launches_per_select(batch.n_obj, with_plane=kw.get("use_ground", True))
        self.streams, self.graphs, self.outs = [], [], []
        self._next = 0
        for _ in range(self.depth):
            st = torch.cuda.Stream(dev)
            st.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(st):
                for _ in range(max(int(warmup), 1)):     # first calls size the cached workspaces: nothing allocates outside the pool later
                    select(batch, n_prop, **kw)
            torch.cuda.synchronize(dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                out = select(batch, n_prop, **kw)
            self.streams.append(st); self.graphs.append(g); self.outs.append(out)
        self.stream, self.graph, self.out = self.streams[0], self.graphs[0], self.outs[0]

    def replay(self):
        if self.depth == 1:
            self.graph.replay()
            return self.out
        k = self._next
        self._next = (k + 1) % self.depth
        st = self.streams[k]
        st.wait_stream(torch.cuda.current_stream(self.batch.device))
        with torch.cuda.stream(st):
            self.graphs[k].replay()
        return self.outs[k]

    def join(self):
        """the caller's stream waits for every replay issued so far (no-op at depth 1: those run on the caller's stream)"""
        if self.depth > 1:
            cur = torch.cuda.current_stream(self.batch.device)
            for st in self.streams:
                cur.wait_stream(st)


class Pipeline:
    """Streams pinned-host batches through one GPU: chunks of `chunk_frames` frames, two device slots, a copy
    stream and a compute stream.  The result (winners [N, 15+4+1+1] packed as fp32) is read back per chunk."""

    def __init__(self, H, W, n_prop, chunk_frames=32, max_obj_per_chunk=None, device="cuda", seed=0,
                 use_ground=True, packed_masks=False, host_pack=False, pack_threads=None):
        self.H, self.W, self.n_prop = H, W, n_prop
        self.dev = torch.device(device)
        self.chunk = chunk_frames
        self.max_obj = max_obj_per_chunk or chunk_frames * 40
        self.seed = seed
        self.use_ground = use_ground
        # host_pack: the host batch holds bool / uint8 masks (as the reference's callers do) and every chunk's masks are
        # packed to 1 bit/px on the host cores (tdod_host_pack_masks) into a pinned staging buffer while the previous
        # chunk's copy is in flight: 1/8 of the mask bytes cross PCIe, K-B reads the bitmap as it is
        self.host_pack = bool(host_pack)
        packed_masks = bool(packed_masks) or self.host_pack
        self.packed_masks = packed_masks      # the device side reads masks at 1 bit/px (`masks_packed` of the host batch, or packed here)
        if pack_threads is None:
            pack_threads = max(1, min(32, (os.cpu_count() or 1) // max(1, int(os.environ.get("WORLD_SIZE", "1")))))
        self.pack_threads = int(pack_threads)
        self.copy_stream = torch.cuda.Stream(self.dev)
        self.compute_stream = torch.cuda.Stream(self.dev)
        d = self.dev
        self.slots = []
        for _ in range(2):
            self.slots.append(dict(
                depth=torch.empty((chunk_frames, H, W), dtype=f32, device=d),
                ground=torch.empty((chunk_frames, H, W), dtype=u8, device=d),
                K=torch.empty((chunk_frames, 3, 3), dtype=f32, device=d),
                boxes=torch.empty((self.max_obj, 4), dtype=f32, device=d),
                pm=torch.empty((self.max_obj, 3), dtype=f32, device=d),
                ps=torch.empty((self.max_obj, 3), dtype=f32, device=d),
                of=torch.empty((self.max_obj,), dtype=i32, device=d),
                masks=None if packed_masks else torch.empty((self.max_obj, H, W), dtype=u8, device=d),
                mpk=torch.empty((self.max_obj, H, (W + 31) // 32), dtype=i32, device=d) if packed_masks else None,
                packed=torch.empty((self.max_obj, 21), dtype=f32, device=d),
                stage=torch.empty((self.max_obj, H, (W + 31) // 32), dtype=i32).pin_memory() if self.host_pack else None,
                copied=torch.cuda.Event(), done=torch.cuda.Event()))
        self.launches = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def run_host(self, hb, obj_off, out=None, obj_id_base=0, frame_id_base=0):
        """hb: FrameBatch of PINNED host tensors; obj_off: python list [F+1] (CSR).  Returns a pinned host
        tensor [N,21]: cube(15), box(4), score, index (as fp32; exact below 2^24).  Synchronises at the end."""
        F, N = hb.n_frames, hb.n_obj
        if out is None:
            out = torch.empty((N, 21), dtype=f32).pin_memory()
        if self.host_pack:
            if hb.masks is None:
                raise ValueError("host_pack: the host batch needs masks (bool / uint8)")
            masks_h = hb.masks.view(u8) if hb.masks.dtype == torch.bool else hb.masks
            if not masks_h.is_contiguous() or masks_h.dtype != u8:
                raise ValueError("host_pack: masks must be a contiguous bool / uint8 tensor")
        elif self.packed_masks:
            if hb.masks_packed is None:
                raise ValueError("this Pipeline was built for packed masks: the host batch needs masks_packed")
            masks_h = hb.masks_packed
        else:
            masks_h = hb.masks.view(u8) if hb.masks.dtype == torch.bool else hb.masks
        ground_h = None
        if hb.ground is not None:
            ground_h = hb.ground.view(u8) if hb.ground.dtype == torch.bool else hb.ground
        k = 0
        for f0 in range(0, F, self.chunk):
            f1 = min(F, f0 + self.chunk)
            o0, o1 = obj_off[f0], obj_off[f1]
            nf, no = f1 - f0, o1 - o0
            if no > self.max_obj:
                raise ValueError("chunk holds %d objects, more than max_obj_per_chunk=%d" % (no, self.max_obj))
            sl = self.slots[k & 1]
            k += 1
            if self.host_pack and no:
                sl["copied"].synchronize()                     # the staging buffer's previous copy has left the host
                L.check(L.lib().tdod_host_pack_masks(masks_h[o0:o1].data_ptr(), no, self.H, self.W, sl["stage"].data_ptr(),
                                                     self.pack_threads))
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(sl["done"])       # slot free again (previous use finished)
                sl["depth"][:nf].copy_(hb.depth[f0:f1], non_blocking=True)
                if ground_h is not None:
                    sl["ground"][:nf].copy_(ground_h[f0:f1], non_blocking=True)
                sl["K"][:nf].copy_(hb.K[f0:f1], non_blocking=True)
                if no:
                    sl["boxes"][:no].copy_(hb.boxes[o0:o1], non_blocking=True)
                    sl["pm"][:no].copy_(hb.prior_mean[o0:o1], non_blocking=True)
                    sl["ps"][:no].copy_(hb.prior_std[o0:o1], non_blocking=True)
                    sl["of"][:no].copy_(hb.obj_frame[o0:o1], non_blocking=True)
                    if self.host_pack:
                        sl["mpk"][:no].copy_(sl["stage"][:no], non_blocking=True)
                    else:
                        sl["mpk" if self.packed_masks else "masks"][:no].copy_(masks_h[o0:o1], non_blocking=True)
                sl["copied"].record(self.copy_stream)
            mask_bytes = self.H * ((self.W + 31) // 32) * 4 if self.packed_masks else self.H * self.W
            self.h2d_bytes += nf * self.H * self.W * (5 if ground_h is not None else 4) + nf * 36 + no * (44 + mask_bytes)
            if no == 0:
                continue
            with torch.cuda.stream(self.compute_stream):
                self.compute_stream.wait_event(sl["copied"])
                b = FrameBatch(sl["depth"][:nf], sl["ground"][:nf] if ground_h is not None else None, sl["K"][:nf],
                               sl["boxes"][:no], sl["pm"][:no], sl["ps"][:no], sl["of"][:no] - f0,
                               None if self.packed_masks else sl["masks"][:no],
                               masks_packed=sl["mpk"][:no] if self.packed_masks else None)
                s = select(b, self.n_prop, seed=self.seed, use_ground=self.use_ground,
                           obj_id_base=obj_id_base + o0, frame_id_base=frame_id_base + f0)
                p = sl["packed"][:no]
                p[:, 0:15] = s.best_cube
                p[:, 15:19] = s.best_box
                p[:, 19] = s.best_score
                p[:, 20] = s.best_idx.to(f32)
                out[o0:o1].copy_(p, non_blocking=True)
                sl["done"].record(self.compute_stream)
            self.launches += LAUNCHES_PER_SELECT if self.use_ground else LAUNCHES_PER_SELECT_NO_GROUND
            self.d2h_bytes += no * 21 * 4
        self.compute_stream.synchronize()
        return out
