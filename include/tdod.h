/* libtdod -- C ABI of the B200-native (sm_100a) proposal-and-scoring path of luchsonice/3dod.
 *
 * The reference has no FFI on this path: it is a Python module API.  Each entry
 * point below names the reference interface it replaces (file:line under /root/reference); the Python
 * mirror in 3dod_b200/ProposalNetwork/ binds them with ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host; the caller allocates inputs and
 *     outputs and keeps them alive until `stream` has drained; the library allocates nothing persistent
 *   - fp32 row-major contiguous unless stated; masks are 1 byte per pixel (torch bool / uint8)
 *   - `stream` is a cudaStream_t passed as void*; no entry point synchronises or copies to the host
 *   - return value: 0 ok; < 0 argument / shape error (checked before any launch, see tdod_strerror);
 *     > 0 a cudaError_t from the launch
 *   - re-entrant and stateless: all state is in the arguments
 */
#ifndef TDOD_H
#define TDOD_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TDOD_OK 0
#define TDOD_EINVAL (-1)   /* null pointer / non-positive size */
#define TDOD_ERANGE (-2)   /* size outside what the kernels support (see message) */

#define TDOD_RNG_HOST 0    /* parity mode: standard normals + yaw ids drawn by the caller */
#define TDOD_RNG_PHILOX 1  /* throughput mode: in-kernel Philox4x32-10 */

/* where the cube centres come from (tdod_pss_args.centre_mode) */
#define TDOD_CENTRE_DEPTH 0    /* proposals.py:350-396 (propose, propose_random_dim, propose_aspect_ratio,
                                  propose_random_rotation): depth gather + pinhole trig seeds, median / std, normal draws */
#define TDOD_CENTRE_UNIFORM 1  /* proposals.py:22-25 (propose_random): x = u*4-2, y = u*2-1, z = u*4+1 */
#define TDOD_CENTRE_PATCH 2    /* proposals.py:51-64 (propose_xy_patch): x, y = vectorized_linspace over the inner half
                                  of the box in normalised image space (conversions.py:50-67), z = u*4+1 */
#define TDOD_CENTRE_PATCH_Z 3  /* proposals.py:98-115 (propose_z): x, y as PATCH; z handed in by the caller (the
                                  per-object torch.quantile / torch.linspace ramp), noise channel 5 taken as is */
/* where the cube dimensions come from (tdod_pss_args.dims_mode) */
#define TDOD_DIMS_PRIOR 0      /* proposals.py:377-383: truncated normals around the class priors */
#define TDOD_DIMS_UNIFORM 1    /* proposals.py:28-30 rescale_interval(u, 0.05, 2) for w, h, l */
#define TDOD_DIMS_ASPECT 2     /* proposals.py:230-239 (propose_aspect_ratio): w uniform, h = w*aux[0], l = w*aux[1] */

/* how torch.std of the centre seeds (proposals.py:386-396) is rounded (tdod_pss_args.std_mode) */
#define TDOD_STD_EXACT 0       /* two-pass fp64, correctly rounded: what torch computes on CPU tensors */
#define TDOD_STD_TORCH_CUDA 1  /* torch's CUDA kernel: fp32 Welford merged in the order of its launch configuration for the
                                  frame's [N_objects_of_frame, n_prop] tensor -- the reference as it runs on the GPU */

/* mode B: the scores the MABO branch adds (roi_heads.py:535-538), computed for all objects by the same call
 * (tdod_pss_args.mode_flags) */
#define TDOD_MODEB_SEG 1       /* score_segmentation + score_mod_segmentation (scorefunction.py:88-126) -> all_scores channels 4, 5 */
#define TDOD_MODEB_PC 2        /* score_point_cloud (scorefunction.py:9-43) -> all_pc */

#define TDOD_BACKPROJECT_NUMPY 0 /* roi_heads.py:346-371: fp64, full-res intrinsics on the strided grid */
#define TDOD_BACKPROJECT_TORCH 1 /* roi_heads.py:1084-1112: fp32, K // stride, centre of the strided grid */

int tdod_version(void);
const char* tdod_strerror(int code);
/* last argument-check message of the calling thread (empty string if none) */
const char* tdod_last_error_string(void);

/* ---- geometry ------------------------------------------------------------------------------------------ */

/* Cubes.get_all_corners (ProposalNetwork/utils/spaces.py:192-204) ->
 * cubercnn/util/math_util.py:142-245 get_cuboid_verts_faces.  cubes [M,15] -> out [M,8,3]. */
int tdod_cube_corners(const float* cubes, int64_t M, float* out_corners3d, void* stream);

/* Cubes.get_bube_corners(K, clamp=(W,H)) (spaces.py:224-245) and cubes_to_box (conversions.py:25-48).
 * cubes [M,15], K [9]; out_corners2d [M,8,2] and out_boxes [M,4] may each be NULL. */
int tdod_project_cubes(const float* cubes, int64_t M, const float* K, int clamp_w, int clamp_h,
                       float* out_corners2d, float* out_boxes, void* stream);

/* The same for cubes of DIFFERENT frames in one launch: K [F,9], cube_frame [M] int32 picks the intrinsics of each cube.
 * Replaces the per-object loop of the MABO branch, cubes_to_box(gt_cube, K, im_shape) once per object
 * (cubercnn/modeling/roi_heads/roi_heads.py:532), when a batch holds the objects of many frames. */
int tdod_project_cubes_indexed(const float* cubes, int64_t M, const float* K, const int32_t* cube_frame, int clamp_w,
                               int clamp_h, float* out_corners2d, float* out_boxes, void* stream);

/* ---- scores (ProposalNetwork/scoring/scorefunction.py) ------------------------------------------------ */

/* score_iou (scorefunction.py:47-49 -> utils.py:186-192 -> detectron2 pairwise_iou). */
int tdod_score_iou(const float* ref_box /*[4]*/, const float* boxes /*[n,4]*/, int64_t n, float* out, void* stream);

/* score_dimensions (scorefunction.py:144-160).  dims has row stride `dims_stride` floats (3 for a packed
 * [n,3], 15 when pointing at column 3 of a cubes tensor). */
int tdod_score_dimensions(const float* prior_mean, const float* prior_std, const float* dims, int64_t dims_stride,
                          const float* ref_box, const float* boxes, int64_t n, float* out, void* stream);

/* first half of score_corners (scorefunction.py:59-68): per mask, the largest external contour's minimum-area
 * rectangle (cv2.findContours RETR_EXTERNAL -> max contourArea -> minAreaRect -> boxPoints).
 * masks [n_masks,H,W] 1 B/px; out_rects [n_masks,4,2]; out_has_rect [n_masks]: 1 = rectangle written, 0 = empty mask (the
 * caller applies the fallback rectangle of :69-75).  The masks are read as aligned 16-byte chunks: up to 15 bytes before
 * the first and after the last mask byte are touched (never interpreted); torch allocations always cover them.
 * Masks up to 1024 px wide whose run tables fit in shared memory take the fused path (bits never leave the SM);
 * larger ones are packed to a 1 bit/px workspace copy first.  Bytes other than 0 / 1 are valid (non-zero = set) and
 * take the general path.  workspace: tdod_mask_rect_workspace_bytes(n_masks,H,W) bytes of device scratch. */
int64_t tdod_mask_rect_workspace_bytes(int64_t n_masks, int H, int W);
int tdod_mask_rect(const uint8_t* masks, int64_t n_masks, int H, int W, float* out_rects, uint8_t* out_has_rect,
                   void* workspace, int64_t workspace_bytes, void* stream);

/* The two halves of tdod_mask_rect as separate entry points, for callers that want to time them or reuse the packed
 * masks: tdod_mask_pack streams the 1 B/px masks once from HBM into the workspace's 1 bit/px bitmap (+ the occupied
 * row range of every mask); tdod_mask_rect_packed computes the rectangles from a workspace tdod_mask_pack filled for
 * the same (n_masks, H, W).  tdod_mask_rect == pack followed by rect_packed. */
int tdod_mask_pack(const uint8_t* masks, int64_t n_masks, int H, int W, void* workspace, int64_t workspace_bytes,
                   void* stream);
int tdod_mask_rect_packed(int64_t n_masks, int H, int W, float* out_rects, uint8_t* out_has_rect, void* workspace,
                          int64_t workspace_bytes, void* stream);

/* The same rectangles from masks the caller already holds at 1 bit per pixel: bitmap [n_masks,H,Wp] 32-bit words, Wp =
 * ceil(W/32), bit (c & 31) of word (c >> 5) of a row = pixel c, unused high bits of a row's last word zero.  For hosts
 * that store or ship masks packed (1/8 of the PCIe bytes of torch.bool masks); same workspace query. */
int tdod_mask_rect_bitmap(const uint32_t* bitmap, int64_t n_masks, int H, int W, float* out_rects, uint8_t* out_has_rect,
                          void* workspace, int64_t workspace_bytes, void* stream);

/* HOST function (no GPU work): packs masks held in host memory at 1 byte per pixel (torch.bool storage; any non-zero byte
 * is "set") into the bitmap layout above, out_bitmap [n_masks,H,Wp] in host memory, with n_threads host threads.  The
 * transfer codec of the end-to-end path: the masks are 75 % of a step's host->device bytes
 * (cubercnn/modeling/roi_heads/roi_heads.py:432 holds them as bool tensors); packed, they cross PCIe at 1/8 of the size
 * and tdod_mask_rect_bitmap consumes them as they are.  A data-format step on the host side of the boundary, not a CPU
 * implementation of any kernel. */
int tdod_host_pack_masks(const uint8_t* masks, int64_t n_masks, int H, int W, uint32_t* out_bitmap, int n_threads);

/* second half of score_corners (scorefunction.py:69-85 + modified_chamfer_distance :51-56).
 * rect [4,2] (ignored when has_rect == 0: the fallback rectangle of :69-75 is computed from the corners). */
int tdod_score_corners(const float* rect, int has_rect, const float* corners2d /*[n,8,2]*/, int64_t n, float* out,
                       void* stream);

/* score_segmentation (scorefunction.py:88-106, mask_iou utils.py:230-239) when power5 == 0;
 * score_mod_segmentation (:108-126, mod_mask_iou utils.py:241-250, int64 inter**5 wrap) when power5 != 0.
 * mask [H,W] 1 B/px, corners2d [n,8,2]. */
int tdod_score_segmentation(const uint8_t* mask, int H, int W, const float* corners2d, int64_t n, int power5,
                            float* out, void* stream);

/* score_point_cloud (scorefunction.py:9-43).  points [P,3] fp32 (points_f64 == 0) or fp64; cubes [n,15];
 * out int64 [n]. */
int tdod_score_point_cloud(const void* points, int points_f64, int64_t P, const float* cubes, int64_t n,
                           int64_t* out, void* stream);

/* iou_3d (ProposalNetwork/utils/utils.py:194-210 -> pytorch3d.ops.box3d_overlap; used by the MABO branch,
 * cubercnn/modeling/roi_heads/roi_heads.py:518,526): exact IoU3D of oriented boxes given by their 8 corners in the
 * vertex order of cubercnn/util/math_util.py:178-207.  gt_corners [G,8,3], corners [G,n,8,3] -> out_iou [G,n]
 * (box g against its n proposals); out_vol [G,n] intersection volumes or NULL. */
int tdod_box3d_iou(const float* gt_corners, const float* corners, int64_t G, int64_t n, float* out_iou, float* out_vol,
                   void* stream);

/* K-D, segmented order: per object the indices of its k best scores in descending order -- np.argsort(scores)[::-1][:k]
 * (cubercnn/modeling/roi_heads/roi_heads.py:278; k = 10 is the top-10 refinement of :585-586, k = n the full order).
 * NaN first (np.argsort sorts NaN last, [::-1] reverses), equal scores by descending index (NumPy leaves tie order
 * unspecified).  scores [S,n] -> out_idx [S,k] int32, out_val [S,k] or NULL.  n <= 16384. */
int tdod_segmented_order(const float* scores, int64_t S, int n, int k, int32_t* out_idx, float* out_val, void* stream);

/* ROIHeads_Boxer.accumulate_scores (roi_heads.py:277-281) for C score channels per object sharing that object's IoU3D:
 * out_curves[s,c,:n_out] = np.maximum.accumulate(iou3d[s][argsort(scores[s,c])[::-1]])[:n_out].  scores are fp64 (NumPy
 * promotes the products with the int64 point-cloud score to float64, :556-579); n_out = n for the eight MABO curves
 * (:543-550), 1 for the 26 score combinations (:553-579, which keep element [0]). */
int tdod_accumulate_scores(const double* scores /*[S,C,n]*/, const float* iou3d /*[S,n]*/, int64_t S, int C, int n, int n_out,
                           float* out_curves /*[S,C,n_out]*/, void* stream);

/* ---- plane ---------------------------------------------------------------------------------------------- */

/* Depth -> strided point cloud (roi_heads.py:346-371 / :1084-1112), row-major compaction.
 * depth [F,H,W]; ground [F,H,W] 1 B/px or NULL; K [F,9].  Outputs per frame f, SoA, capacity
 * P = ceil(H/stride)*ceil(W/stride): plane_pts [F,3,P] (ground>0 points, or all when ground NULL),
 * other_pts [F,3,P] or NULL (ground==0 points), counts [2,F] int32 (row 0: plane points per frame --
 * directly usable as tdod_ransac_plane's n_pts --, row 1: the others).  fp64 when flavour == NUMPY else fp32. */
int tdod_backproject(const float* depth, const uint8_t* ground, const float* K, int F, int H, int W, int stride,
                     int flavour, void* plane_pts, void* other_pts, int32_t* counts, void* stream);

/* Plane.fit_parallel (ProposalNetwork/utils/plane.py:79-134; fp32) / Plane_np.fit == pyransac3d.Plane.fit
 * (plane.py:154-209, roi_heads.py:374-376; fp64), batched over F frames of SoA points [F,3,P_cap].
 * sample_ids [F,iters,3] int32 (python random.sample stream) or NULL -> Philox ids keyed by (`seed`,
 * frame_id_base + f, iteration), so a frame's plane does not depend on how frames are batched or sharded.
 * Outputs: out_eq [F,4] (same dtype as the points; sign as the NumPy flavour, i.e. NOT negated),
 * out_best_iter [F], out_count [F], out_normal [F,3] rectified up-vector (roi_heads.py:411-428), fp64/fp32,
 * out_inlier_mask [F,P_cap] uint8 or NULL.  out_normal is rectified from -eq for fp32 points (fit_parallel
 * returns the negated equation, plane.py:134) and from +eq for fp64 points (pyransac3d convention).
 * workspace: tdod_ransac_workspace_bytes(F, iters, pts_f64) bytes of device scratch. */
int64_t tdod_ransac_workspace_bytes(int F, int iters, int pts_f64);
int tdod_ransac_plane(const void* pts, int pts_f64, const int32_t* n_pts, int F, int64_t P_cap,
                      const int32_t* sample_ids, uint64_t seed, int64_t frame_id_base, int iters, double thresh,
                      void* out_eq, int32_t* out_best_iter, int32_t* out_count, void* out_normal, uint8_t* out_inlier_mask,
                      void* workspace, int64_t workspace_bytes, void* stream);

/* tdod_ransac_plane with NO workspace argument and NO allocation of any kind (no cudaMalloc / cudaFree, no allocating
 * library call): its scratch is a fixed device array inside the library (2 887 680 bytes, held by every GPU that loads
 * libtdod until the process exits).  This entry point has no counterpart in the reference; it was added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same arguments, outputs and results as tdod_ransac_plane.
 * Limits: F <= TDOD_RANSAC_STATIC_MAX_FRAMES and F * iters <= TDOD_RANSAC_STATIC_MAX_HYPS (e.g. 64 frames x 1000
 * iterations); beyond them the call returns TDOD_ERANGE and launches nothing.
 * Sharing: every call on a device uses the same array.  Calls that overlap on the GPU (different streams without an
 * ordering between them) race; calls ordered on one stream, or by events, are safe.  The array keeps the previous call's
 * contents; the part that must start at zero (the inlier counts) is cleared by the call itself on its stream.
 * The first call on a device resolves the array's device address (cudaGetSymbolAddress, that device being current):
 * make that first call outside a stream capture. */
#define TDOD_RANSAC_STATIC_MAX_HYPS 65536
#define TDOD_RANSAC_STATIC_MAX_FRAMES 1024
int tdod_ransac_plane_static(const void* pts, int pts_f64, const int32_t* n_pts, int F, int64_t P_cap,
                             const int32_t* sample_ids, uint64_t seed, int64_t frame_id_base, int iters, double thresh,
                             void* out_eq, int32_t* out_best_iter, int32_t* out_count, void* out_normal,
                             uint8_t* out_inlier_mask, void* stream);

/* orthobasis_from_normal_t over linspace(0, pi, 36) (utils.py:112-146, proposals.py:405-406).
 * normals [F,3] fp64 (normals_f64) or fp32 -> out_table [F,36,9] fp32 (row-major 3x3, columns x, n, n x x). */
int tdod_yaw_table(const void* normals, int normals_f64, int F, float* out_table, void* stream);

/* ---- ingest ---------------------------------------------------------------------------------------------- */

/* Depth / ground map resize to the network resolution, PIL Image.resize(.., NEAREST) semantics
 * (cubercnn/data/dataset_mapper.py:127-128, :141-142).  src [F,h,w] of 1- or 4-byte elements -> dst [F,H,W];
 * xmap [W], ymap [H] int32 source indices (Pillow's accumulating double-precision stepping, built by the caller). */
int tdod_resize_nearest(const void* src, int elem_bytes, int F, int h, int w, const int32_t* xmap, const int32_t* ymap,
                        void* dst, int H, int W, void* stream);

/* torch.std(x, dim=1) of a contiguous fp32 tensor x [N,n] exactly as torch's CUDA kernel rounds it (the TDOD_STD_TORCH_CUDA
 * flavour above, exposed for checking it against torch itself).  out [N]. */
int tdod_torch_std(const float* x, int64_t N, int n, float* out, void* stream);

/* ---- fused propose + project + score + select (the hot kernel) --------------------------------------- */

typedef struct tdod_pss_args {
    /* frames */
    const float* depth;        /* [F,H,W] */
    const float* K;            /* [F,9] */
    const float* yaw_table;    /* [F,36,9] or NULL -> random rotations (proposals.py:401-402) */
    int32_t F, H, W;
    /* objects */
    int32_t n_obj, n_prop;
    const int32_t* obj_frame;  /* [n_obj] */
    const float* boxes;        /* [n_obj,4] xyxy: sampling range + aspect-ratio reference */
    const float* iou_boxes;    /* [n_obj,4] or NULL (= boxes): IoU reference (roi_heads.py:466 vs :496) */
    const float* prior_mean;   /* [n_obj,3] (w,h,l) */
    const float* prior_std;    /* [n_obj,3] */
    const float* rects;        /* [n_obj,4,2] from tdod_mask_rect */
    const uint8_t* has_rect;   /* [n_obj] or NULL (all present) */
    /* randomness */
    int32_t rng_mode;          /* TDOD_RNG_HOST | TDOD_RNG_PHILOX */
    const float* noise;        /* HOST mode: [n_obj,6,n_prop], order w,h,l,x,y,z: accepted standard normals, or U[0,1)
                                  draws for the channels an ablation mode samples uniformly (unused channels ignored) */
    const int16_t* yaw_idx;    /* HOST mode with yaw_table: [n_obj,n_prop] in [0,36) */
    const float* rot;          /* HOST mode without yaw_table: [n_obj,n_prop,9] */
    uint64_t seed;             /* PHILOX mode */
    int64_t obj_id_base;       /* PHILOX mode: global id of object 0 (results independent of sharding) */
    /* outputs: winners (always) */
    int32_t* best_idx;         /* [n_obj] */
    float* best_score;         /* [n_obj] */
    float* best_cube;          /* [n_obj,15] */
    float* best_box;           /* [n_obj,4] */
    /* outputs: everything (each may be NULL) */
    float* all_cubes;          /* [n_obj,n_prop,15] */
    float* all_boxes;          /* [n_obj,n_prop,4] */
    float* all_scores;         /* [n_obj,score_channels,n_prop]: iou, dim, corners, combined (, segmentation, mod-segmentation) */
    float* stats;              /* [n_obj,6]: mean/std of the x, y, z samplers */
    int32_t cluster_size;      /* CTAs per object: 0 = auto, else 1,2,4,8 */
    int32_t propose_only;      /* != 0: sample + project only (proposals.propose / cubes_to_box): fills all_cubes /
                                  all_boxes / stats, needs no rects and writes no winners */
    /* ablation samplers (proposals.py:20-336); all zero = the production sampler `propose` */
    int32_t centre_mode;       /* TDOD_CENTRE_* */
    int32_t dims_mode;         /* TDOD_DIMS_* */
    int32_t im_w, im_h;        /* im_shape of pixel_to_normalised_space (PATCH modes); 0 -> W, H */
    const float* aux;          /* [n_obj,2]: DIMS_ASPECT (ratio_h, ratio_l); NULL otherwise */
    int32_t std_mode;          /* TDOD_STD_* */
    /* mode B (all zero / NULL = mode A only).  Needs all_cubes (in this mode every cube is an output of the call: the caller
     * scores them against the ground truth, roi_heads.py:526); corners, hulls and polygons are never materialised */
    int32_t mode_flags;        /* TDOD_MODEB_* */
    int32_t score_channels;    /* channels of all_scores: 0 or 4 = iou, dim, corners, combined; 6 = + segmentation, mod-segmentation */
    const uint8_t* masks;      /* [n_obj,H,W] 1 B/px (TDOD_MODEB_SEG) */
    const void* points;        /* [F,3,P_cap] SoA, the frames' non-ground points as tdod_backproject writes them (TDOD_MODEB_PC) */
    const int32_t* n_points;   /* [F] */
    int64_t P_cap;
    int32_t points_f64;        /* points are fp64 (NumPy flavour, roi_heads.py:535) or fp32 */
    int32_t* all_pc;           /* [n_obj,n_prop] points strictly inside each proposal's intervals (int64 in the reference) */
    /* replay / audit outputs (each may be NULL): the numbers behind the proposal set, in HOST-mode layout, so that a PHILOX
     * run can be fed back through HOST mode (and through the oracle's cubes_from_noise) bit for bit */
    float* out_noise;          /* [n_obj,6,n_prop]: accepted standard normals (or U[0,1) draws) for w,h,l,x,y,z */
    int16_t* out_yaw_idx;      /* [n_obj,n_prop]: yaw bin (with yaw_table) */
    float* out_seeds;          /* [n_obj,3,n_prop]: the centre seeds x, y, z_tmp + l/2 (proposals.py:366-375, :394) whose
                                  median / std parameterise the centre draws */
} tdod_pss_args;

/* propose (proposals.py:338-424) + cubes_to_box + score_iou + score_dimensions + score_corners +
 * combine/argmax (roi_heads.py:493-505) in one launch; one CTA cluster per object. */
int tdod_propose_score_select(const tdod_pss_args* args_host, void* stream);

/* ---- weak-loss geometry of ROIHeads3DScore -------------------------------------------- */

/* Forward of the loop roi_heads.py:1571-1583 (+ :1599) for N objects, each with its own K [N,9] and image size
 * im_wh [N,2] int32 (width, height): Cubes.get_bube_corners (spaces.py:224-245) -> out_corners [N,8,2] (clamped to
 * the wide range of :241-242); the second clamp to the image (roi_heads.py:1574-1575) -> out_seg_corners [N,8,2] or
 * NULL; cubes_to_box (conversions.py:25-48) -> out_boxes [N,4]; out_arg [N,4] int32 = corner index of each box
 * coordinate (torch.min / torch.max along the corner dim: first occurrence, NaN first) for the backward.
 * gt_boxes [N,4] or NULL: torchvision.ops.generalized_box_iou_loss(gt, box, reduction='none') -> out_giou [N] and its
 * derivative with respect to the box out_dgiou [N,4]. */
int tdod_cube_box_fwd(const float* cubes, const float* K, const int32_t* im_wh, const float* gt_boxes, int64_t N,
                      float* out_corners, float* out_seg_corners, float* out_boxes, int32_t* out_arg, float* out_giou,
                      float* out_dgiou, void* stream);

/* Backward of tdod_cube_box_fwd (what torch.autograd builds for the loop above): incoming gradients with respect to
 * out_corners, out_seg_corners, out_boxes [N,4], out_giou [N] (each may be NULL) -> out_g_cubes [N,15]
 * (centre, w h l, R row-major).  clamp passes the gradient on the closed interval, min / max to the index in `arg`. */
int tdod_cube_box_bwd(const float* cubes, const float* K, const int32_t* im_wh, const int32_t* arg, int64_t N,
                      const float* g_corners, const float* g_seg_corners, const float* g_boxes, const float* g_giou,
                      const float* dgiou, float* out_g_cubes, void* stream);

/* z_loss (roi_heads.py:1151-1194): per cube the depth is moved in `steps` steps of values[s] (torch.linspace(0, 4.9,
 * 50)) away from / towards the camera until the projected box area is closest to the ground-truth box area;
 * out [N] = |z - z_best| / 2, or 0.1 * steps / 2 when the centre test of :1166-1167 fails (reproduced as written). */
int tdod_z_sweep(const float* gt_boxes, const float* cubes, const float* K, const int32_t* im_wh, const float* proj_boxes,
                 const float* values, int steps, int64_t N, float* out, void* stream);

/* jarvis_march (ProposalNetwork/utils/utils.py:424-468, with findDuplicates :397-422) on n sets of 8 points [n,8,2]:
 * out_hull [n,8,2] hull vertices in the order the reference returns them, out_meta [n,9] int32 = {m, corner index of
 * vertex 0..m-1, -1 padding}. */
int tdod_hull(const float* corners2d, int64_t n, float* out_hull, int32_t* out_meta, void* stream);

/* fill_polygon (utils.py:470-502) + the pixel sums every loss of segment_loss (roi_heads.py:1030-1052) is made of:
 * with t = soft hull mask at a pixel and x = masks[mask_idx[i]] at it (masks [M,H,W] 1 B/px), out_moments [n,12] fp64 =
 * sum t^k over x==0 (k=0..4), sum t^k over x==1 (k=0..4), sum sigmoid(t) over x==0, over x==1 (the last two only
 * when with_sigmoid != 0 -- dice_loss needs them --, else the t == 0 pixels' 1/2 alone).  The H x W soft mask is never
 * written. */
int tdod_hull_fill_moments(const uint8_t* masks, const int32_t* mask_idx, const float* hull, const int32_t* meta,
                           int64_t n, int H, int W, int with_sigmoid, double* out_moments, void* stream);

/* Backward of the sums above to the 8 corners: g_moments [n,12] fp64 -> out_g_corners [n,8,2] fp32.
 * workspace: n * 24 doubles of device scratch. */
int tdod_hull_fill_grad(const uint8_t* masks, const int32_t* mask_idx, const float* hull, const int32_t* meta, int64_t n,
                        int H, int W, const double* g_moments, double* workspace, float* out_g_corners, void* stream);

/* convex_hull(mask, coords) itself (utils.py:504-507): the dense soft mask out_mask [n,H,W] fp32 of n hulls, and the
 * backward of sum(g_mask * out_mask) to the 8 corners (workspace: n * 24 doubles). */
int tdod_hull_fill(const float* hull, const int32_t* meta, int64_t n, int H, int W, float* out_mask, void* stream);
int tdod_hull_fill_dense_grad(const float* hull, const int32_t* meta, int64_t n, int H, int W, const float* g_mask,
                              double* workspace, float* out_g_corners, void* stream);

/* Depth order statistics of the pseudo-ground-truth losses, one value per box (n boxes, box i in frame frame[i] of
 * depth [F,H,W] fp32, boxes [n,4] fp32 x1 y1 x2 y2):
 *   mode 0  pseudo_gt_z_box_loss (roi_heads.py:1196-1231): torch.median of the depth crop under the box clipped to the
 *           image; out_flag = 1.  Boxes whose clipped area is not positive: the depth at their clamped centre, out_flag = 0
 *           (the caller orders the values as the reference does, medians before centre samples within a frame).
 *   mode 1  pseudo_gt_z_point_loss (:1256-1277): depth at (clamp(x1,10,W-11), clamp(y1,10,H-11)); x2, y2 unused.
 *   mode 2  depth_range_loss (:1279-1305): torch.quantile(.,0.9) - torch.quantile(.,0.1) of the depth under
 *           masks[mask_idx[i]] ([M,H,W] uint8, out_flag = 1), of the crop under the box when the mask is empty (0).
 * An empty crop (the reference raises) and a crop that holds a NaN give NaN.  out_flag may be NULL. */
int tdod_depth_select(const float* depth, int64_t F, int H, int W, const int32_t* frame, const float* boxes,
                      const uint8_t* masks, const int32_t* mask_idx, int mode, int64_t n, float* out, int32_t* out_flag,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TDOD_H */
