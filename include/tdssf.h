/*
 * tdssf.h — C ABI of the B200-native (sm_100a) cross-modal fusion hot path of 3DSSF.
 *
 * This is the drop-in boundary.  The reference (CharlesChang012/3D-Semantic-Segmentation-Fusion)
 * has no FFI layer of its own: its hot path is a set of Python functions built from ATen ops.
 * Every entry point below names the reference interface it replaces (file:line relative to the
 * reference tree).  The Python mirror of the reference modules (package `utils/` under
 * `3d-semantic-segmentation-fusion_b200/`) binds these with ctypes — see INTEGRATION.md.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the parameter name ends in `_host`;
 *   - `stream` is a `cudaStream_t` passed as `void*` (NULL = legacy default stream);
 *   - nothing is allocated behind the caller's back: outputs and workspaces are caller-owned,
 *     workspace sizes come from the matching `*_workspace_bytes()` call;
 *   - return value 0 = OK, negative = TDSSF_E*; `tdssf_last_error()` gives a thread-local message;
 *   - calls are asynchronous with respect to the host (stream-ordered) unless stated otherwise;
 *   - inputs are never written.
 */
#ifndef TDSSF_H_
#define TDSSF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TDSSF_VERSION 100 /* 0.1.0 */

#if defined(__GNUC__)
#define TDSSF_API __attribute__((visibility("default")))
#else
#define TDSSF_API
#endif

#define TDSSF_OK 0
#define TDSSF_EINVAL (-1)  /* bad argument (shape, alignment, NULL)            */
#define TDSSF_ECUDA (-2)   /* a CUDA runtime call or a launch failed           */
#define TDSSF_ENOSUP (-3)  /* configuration not supported by this build        */
#define TDSSF_EDEVICE (-4) /* the current device is not sm_100 (no fallback)   */

/* token / operand element types */
#define TDSSF_F32 0
#define TDSSF_F16 1
#define TDSSF_BF16 2

/* one 16-byte "camera assignment" record per LiDAR point:
 *   patch[c]  flat patch index (y*grid_w+x, clamped) of the point in camera c, c < 6
 *   mask      bit c set  <=>  the point is valid in camera c (depth>0, inside the image)
 *   count     popcount(mask)
 * This is the compact hand-off between tdssf_project and the gather / fused kernels. */
typedef struct tdssf_rec {
  uint16_t patch[6];
  uint8_t mask;
  uint8_t count;
  uint16_t pad_;
} tdssf_rec;

#define TDSSF_MAX_CAMS 6

TDSSF_API int tdssf_version(void);
TDSSF_API const char* tdssf_last_error(void);
/* 0 if the current CUDA device is compute capability 10.x, TDSSF_EDEVICE otherwise. */
TDSSF_API int tdssf_check_device(void);

/* ---------------------------------------------------------------------------------------------
 * a1+a2+a3  projection, mask, pixel rescale, patch index
 * replaces  multi_camera_projector            utils/fusion_model.py:14-86
 *           scale_pixel_coords                utils/fusion_model.py:89-127
 *           patch index in FeatureFusionModel.forward   utils/fusion_model.py:172-180,188
 *
 * xyz        points of sweep b start at xyz + b*sweep_stride; point v at + v*point_stride (floats);
 *            (x,y,z) are the first three floats.  point_stride==4 with a 16-byte aligned base takes
 *            the float4 path.
 * K          [B,NC,3,3] fp32 row-major          Rt  [B,NC,4,4] fp32 row-major (lidar -> camera)
 * img_hw     [B,2] fp32 (H,W) of the ORIGINAL images
 * Outputs (any may be NULL):
 *   pix      [B,NC,V,2] fp32 (u,v) in original pixels, -1 where invalid   (reference layout)
 *   depth    [B,NC,V]   fp32 camera z, -1 where invalid
 *   valid    [B,NC,V]   u8   0/1
 *   flat_idx [B,NC,V]   i32  clamped flat patch index (needs resize_h/w, patch)
 *   rec      [B,V]      tdssf_rec (NC <= 6)
 * resize_h, resize_w, patch: encoder attributes (utils/camera.py:24-34); ignored when flat_idx and
 * rec are both NULL.
 * fp32 evaluation order is the reference's CPU order: extrinsic rows with separately
 * rounded mul/add, intrinsic rows as an FMA chain, IEEE division. */
TDSSF_API int tdssf_project(const float* xyz, int64_t sweep_stride, int64_t point_stride,
                  const float* K, const float* Rt, const float* img_hw,
                  int B, int V, int NC,
                  int resize_h, int resize_w, int patch,
                  float* pix, float* depth, uint8_t* valid, int32_t* flat_idx, tdssf_rec* rec,
                  void* stream);

/* a2 stand-alone   replaces scale_pixel_coords   utils/fusion_model.py:89-127
 * pix/out [B, per_sweep, 2] fp32 (u,v); img_hw [hw_rows,2] fp32 (H,W) with hw_rows == B or 1 (broadcast);
 * out = pix * (new_w / max(W,1e-6), new_h / max(H,1e-6)), scales computed in fp32 like the reference. */
TDSSF_API int tdssf_scale_pixels(const float* pix, const float* img_hw, int hw_rows, int B, int64_t per_sweep,
                       float new_h, float new_w, float* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a4+a5  per-camera patch-token gather and masked mean over cameras
 * replaces  utils/fusion_model.py:184-201
 * tokens  [B,NC,M,Ci]  fp32 / fp16 / bf16 (tok_dtype), row-major, 16-byte aligned, Ci % 8 == 0
 * rec     [B,V]
 * feat    [B,V,Ci] fp32.  fp16/bf16 tokens: the mean is rounded to the token dtype as the reference's
 *         same-dtype arithmetic does, then widened (the reference's torch.cat promotes to fp32).
 * Precondition: tokens are finite (the reference reads and multiplies by 0 for invalid cameras;
 * this kernel skips the load — SURVEY F4). */
TDSSF_API int tdssf_gather_mean(const void* tokens, int tok_dtype, const tdssf_rec* rec,
                      int B, int V, int NC, int M, int Ci,
                      float* feat, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a6  fusion MLP  448 -> 256 -> 64 -> 16  (concat [point feature | image feature] first)
 * replaces  FeatureFusionModel.mlp + torch.cat   utils/fusion_model.py:144-150,203-204
 * Weights are nn.Linear layout (out,in) row-major fp32: W1[256,448] b1[256] W2[64,256] b2[64]
 * W3[16,64] b3[16].  Column order of W1's `in` axis: 64 point-feature columns, then 384 image.
 */
typedef struct tdssf_mlp_weights {
  const float* W1; const float* b1;
  const float* W2; const float* b2;
  const float* W3; const float* b3;
} tdssf_mlp_weights;

/* fp32 CUDA-core path (1e-5 parity mode).  x_pt [N,64], x_img [N,384], logits [N,16].
 * h1 [N,256] / h2 [N,64] (post-ReLU activations) are written when non-NULL (saved for backward). */
TDSSF_API int tdssf_mlp_fwd_f32(const float* x_pt, const float* x_img, int64_t N,
                      const tdssf_mlp_weights* w_host, float* h1, float* h2, float* logits,
                      void* stream);

/* fp32 backward: given dlogits [N,16] and saved x/h1/h2, accumulates (+=) dW*, db* (fp32, caller
 * zeroes them) and writes dx_pt [N,64] (gradient to the PTv3 feature).  No gradient is
 * produced for the image feature (the reference image branch runs under inference_mode).
 * workspace: tdssf_mlp_bwd_workspace_bytes(N). */
typedef struct tdssf_mlp_grads {
  float* dW1; float* db1; float* dW2; float* db2; float* dW3; float* db3;
} tdssf_mlp_grads;
TDSSF_API size_t tdssf_mlp_bwd_workspace_bytes(int64_t N);
TDSSF_API int tdssf_mlp_bwd_f32(const float* x_pt, const float* x_img, const float* h1, const float* h2,
                      const float* dlogits, int64_t N, const tdssf_mlp_weights* w_host,
                      const tdssf_mlp_grads* g_host, float* dx_pt,
                      void* workspace, size_t workspace_bytes, void* stream);

/* bf16 shadow of the weights in the tensor-core operand layout (prepared once per optimizer step).
 * packed must hold tdssf_mlp_pack_bytes() bytes. */
TDSSF_API size_t tdssf_mlp_pack_bytes(void);
TDSSF_API int tdssf_mlp_pack_weights(const tdssf_mlp_weights* w_host, void* packed, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a1..a6 fused inference: project -> gather -> masked mean -> concat -> MLP (tcgen05, bf16 operands,
 * fp32 accumulate) -> logits, without materialising the 1 536 B/point image feature.
 * replaces  FeatureFusionModel.forward up to voxel_scores   utils/fusion_model.py:152-204
 * x_pt    [B*V,64] fp32 point features (PTv3 output)
 * logits  [B*V,16] fp32
 * rec_out optional [B,V] camera-assignment records (NULL to skip).
 * x_bf16 / h1_bf16 / h2_bf16  optional (NULL to skip; training): [B*V,448] the MLP input rows [point feature | image
 *          feature] and [B*V,256] / [B*V,64] the post-ReLU activations, all in bf16 exactly as the tensor cores consumed
 *          them - what tdssf_mlp_bwd_tc needs, so that training never materialises the fp32 image feature either.
 *          h1_bf16 / h2_bf16 32-byte aligned (256-bit stores; tdssf_mlp_bwd_tc asks the same).
 * preds_out optional [B*V] int64 class ids = argmax(logits) + 1 (first maximum, like torch.argmax; utils/losses.py:66),
 *          for the evaluation loop's confusion matrix (NULL to skip).
 * workspace: tdssf_fusion_workspace_bytes(B, V, tok_dtype) bytes, 256-byte aligned.  Holds, per point, the address of
 *          the token row the gather will read, and for the points seen by several cameras their camera mean (worst case
 *          one token row per point; only the rows of such points are ever touched).  Scratch: nothing persists between
 *          calls, and the same workspace may serve any call with B*V not larger than it was sized for. */
TDSSF_API size_t tdssf_fusion_workspace_bytes(int B, int V, int tok_dtype);
TDSSF_API int tdssf_fusion_fwd(const float* xyz, int64_t sweep_stride, int64_t point_stride,
                     const float* K, const float* Rt, const float* img_hw,
                     int B, int V, int NC, int resize_h, int resize_w, int patch,
                     const void* tokens, int tok_dtype, int M, int Ci,
                     const float* x_pt, const void* packed_weights,
                     float* logits, tdssf_rec* rec_out, int64_t* preds_out,
                     void* x_bf16, void* h1_bf16, void* h2_bf16,
                     void* workspace, size_t workspace_bytes, void* stream);

/* tensor-core MLP on pre-gathered features (training forward, and the unfused check of the fused kernel).
 * x_img fp32 [N,384].  h1_bf16 [N,256] / h2_bf16 [N,64]: optional (NULL to skip) post-ReLU activations in bf16, exactly
 * the values the next layer consumed, saved for tdssf_mlp_bwd_tc. */
TDSSF_API int tdssf_mlp_fwd_tc(const float* x_pt, const float* x_img, int64_t N, const void* packed_weights,
                     float* logits, void* h1_bf16, void* h2_bf16, void* stream);

/* bf16 backward of the MLP from the activations tdssf_mlp_fwd_tc saved (training with the tensor-core forward).
 * replaces  the autograd of FeatureFusionModel.mlp   utils/fusion_model.py:144-150,203-204 (utils/train.py:91)
 * The input rows come either as x_bf16 [N,448] (saved by tdssf_fusion_fwd) or, when x_bf16 is NULL, as fp32 x_pt / x_img.
 * Accumulates (+=) dW*, db* into the caller's fp32 buffers (zeroed by the caller) and writes dx_pt [N,64] (may be NULL).
 * Every product is a hand-written kernel of this library (csrc/mlp_bwd_tc.cu; no vendor GEMM library): the data-gradient
 * chain dl@W3 -> mask -> @W2 -> mask -> @W1[:, :64] per 128-point tile with tcgen05.mma (the ReLU masks, bias sums and bf16
 * copies in its epilogues), and the weight gradients dW1 / dW2 with the points as the K dimension of tcgen05.mma on MN-major
 * operands delivered by TMA, accumulated per CTA in TMEM and reduced in a fixed order (deterministic); dW3 / db3 on the CUDA cores.
 * All pointers 32-byte aligned (256-bit row accesses).  workspace: tdssf_mlp_bwd_tc_workspace_bytes(N), 256-byte aligned. */
TDSSF_API size_t tdssf_mlp_bwd_tc_workspace_bytes(int64_t N);
TDSSF_API int tdssf_mlp_bwd_tc(const float* x_pt, const float* x_img, const void* x_bf16, const void* h1_bf16, const void* h2_bf16,
                     const float* dlogits, int64_t N, const tdssf_mlp_weights* w_host,
                     const tdssf_mlp_grads* g_host, float* dx_pt,
                     void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a7  devoxelize: nearest voxel per raw point, gather, and the backward scatter-add
 * replaces  LiDARFeatureEncoder.devoxelize   utils/lidar.py:120-172
 * pts [B,P,pts_stride] / vox [B,V,vox_stride] (xyz first).  idx [B,P] int64 = argmin_v |p-v|^2,
 * first index wins ties.  All V rows take part, zero padding included.
 * workspace == NULL: brute force over all P x V pairs.  With a workspace of tdssf_devox_nn_workspace_bytes(B,P,V)
 * bytes (256-byte aligned): the voxels are hashed into cubic cells of edge `cell` (same length unit as the
 * coordinates; any positive value gives the same result, a few voxel sizes is fastest) and each point searches the 27
 * cells around it; points without a voxel closer than 0.99*cell fall back to the full scan.  Same indices as the brute
 * force, bit for bit. */
TDSSF_API size_t tdssf_devox_nn_workspace_bytes(int B, int P, int V);
TDSSF_API int tdssf_devox_nn(const float* pts, int64_t pts_stride, const float* vox, int64_t vox_stride,
                   int B, int P, int V, int64_t* idx, float cell, void* workspace, size_t workspace_bytes,
                   void* stream);
/* tdssf_devox_nn (grid-accelerated path) for callers that cannot allocate device memory - code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
 * Same inputs, same indices, but the hash-grid workspace is a static device array of
 * TDSSF_DEVOX_NN_STATIC_BYTES inside the library (one per GPU that loads it, held until exit); the call allocates
 * nothing.  Limit: tdssf_devox_nn_workspace_bytes(B, P, V) <= TDSSF_DEVOX_NN_STATIC_BYTES (16 MiB: e.g. 8 sweeps of
 * 34 720 raw points and voxels), beyond which it returns TDSSF_EINVAL with the sizes in tdssf_last_error().
 * All calls on a device share that array: a call re-initialises what it reads, but two calls that overlap on the GPU
 * (different streams without ordering, or two host threads) race - serialise them, or use tdssf_devox_nn. */
#define TDSSF_DEVOX_NN_STATIC_BYTES (16u << 20)
TDSSF_API int tdssf_devox_nn_static(const float* pts, int64_t pts_stride, const float* vox, int64_t vox_stride,
                                    int B, int P, int V, int64_t* idx, float cell, void* stream);
/* out[b,p,:] = scores[b,idx[b,p],:]   scores [B,V,C], out [B,P,C] */
TDSSF_API int tdssf_devox_gather(const float* scores, const int64_t* idx, int B, int P, int V, int C,
                       float* out, void* stream);
/* dscores[b,idx[b,p],:] += dout[b,p,:]   (dscores zeroed by the caller) */
TDSSF_API int tdssf_devox_scatter_add(const float* dout, const int64_t* idx, int B, int P, int V, int C,
                            float* dscores, void* stream);

/* ---------------------------------------------------------------------------------------------
 * f1  voxelizer: voxel down-sampling + 5-NN weighted intensity + grid coordinates, on the device
 * replaces  LiDARFeatureEncoder.voxelize_open3d   utils/lidar.py:83-118  (Open3D voxel_down_sample on the host,
 *           scikit-learn NearestNeighbors(5), numpy) for B clouds at once.
 * pts [B,P,point_stride_floats] fp32 (xyz first, intensity at column 3 when has_intensity).
 * voxel_raw [B,P,4] fp32: rows 0..n_voxels[b]-1 = (centroid xyz, weighted intensity), the rest zero padding.
 * voxel_coords [B,P,3] int32 = floor(centroid / voxel_size).  n_voxels [B] int32 (device).
 * Arithmetic follows Open3D 0.18 (all in double on the fp32 coordinates: voxel index
 * int(floor((p - (min_bound - voxel_size/2)) / voxel_size)), points accumulated in point order, sum / count) and the
 * reference's numpy (intensity = sum_k I[nbr_k] * {0.4,0.3,0.2,0.05,0.05}[k] in double, nearest first; equidistant
 * neighbours: lower point index first, where scikit-learn's order is unspecified).  Voxels come in the order of their
 * first point (Open3D: unordered_map iteration order).  knn_cell: edge of the neighbour-search grid cells, any
 * positive value gives the same result; <= 0 selects max(2*voxel_size, 0.25).
 * status (optional, device int32): bit 0 set when a coordinate was not finite or an axis spans more than 2^21 voxels
 * (results undefined then).  Errors: TDSSF_EINVAL for voxel_size <= 0 or has_intensity with P < 5 (the reference
 * raises in both cases). */
TDSSF_API size_t tdssf_voxelize_workspace_bytes(int B, int P);
TDSSF_API int tdssf_voxelize(const float* pts, int64_t point_stride_floats, int B, int P, int has_intensity,
                   double voxel_size, double knn_cell, float* voxel_raw, int32_t* voxel_coords, int32_t* n_voxels,
                   int32_t* status, void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a8+a9+a10  CE + Lovasz-Softmax loss, forward and backward in one call
 * replaces  CELSLoss.forward            utils/losses.py:19-70
 *           lovasz_softmax_flat         LovaszSoftmax/pytorch/lovasz_losses.py:171-199
 *           lovasz_grad                 LovaszSoftmax/pytorch/lovasz_losses.py:18-30
 * logits  [n_rows,16] fp32    labels [n_rows] int64 in 0..16    mask [n_rows] u8 or NULL
 * class_w [16] fp32 (device)
 * Outputs:
 *   out_scalars (device) float[4] = {total, ce, lovasz, n_present}; double-precision partials inside
 *   n_valid_host: number of rows that survive mask & (label != ignore) — written after an internal
 *                 stream synchronisation (the reference's boolean indexing synchronises here too)
 *   preds, gt  [>= n_rows] int64 — first n_valid entries are meaningful, original row order
 *   dlogits    [n_rows,16] fp32 or NULL: d total / d logits (zeros on dropped rows)
 * Sort order of the Lovasz errors: descending value, ties by ascending original index.
 * workspace: tdssf_cels_workspace_bytes(n_rows). */
TDSSF_API size_t tdssf_cels_workspace_bytes(int64_t n_rows);
TDSSF_API int tdssf_cels_fwd_bwd(const float* logits, const int64_t* labels, const uint8_t* mask,
                       int64_t n_rows, const float* class_w, float lambda_lovasz, int ignore_index,
                       float* out_scalars, int64_t* n_valid_host, int64_t* preds, int64_t* gt,
                       float* dlogits, void* workspace, size_t workspace_bytes, void* stream);

/* The same computation without a host synchronisation inside the call, in two stages on one workspace:
 *   tdssf_cels_count            flags + scan only; the number of surviving rows goes to n_valid_dev (device int32), from where
 *                               the caller may start an asynchronous copy to the host
 *   tdssf_cels_fwd_bwd_counted  everything else, with every grid sized for n_rows and the real bounds read from the counts
 *                               in the workspace (same results as tdssf_cels_fwd_bwd).  A call that repeats the arguments of
 *                               a recent one (a training loop) is captured once into a CUDA graph and replayed with one
 *                               launch (TDSSF_CELS_GRAPH=0 disables).  preds / gt hold n_valid meaningful entries. */
TDSSF_API int tdssf_cels_count(const int64_t* labels, const uint8_t* mask, int64_t n_rows, int ignore_index, int32_t* n_valid_dev,
                               void* workspace, size_t workspace_bytes, void* stream);
TDSSF_API int tdssf_cels_fwd_bwd_counted(const float* logits, const int64_t* labels, const uint8_t* mask,
                                         int64_t n_rows, const float* class_w, float lambda_lovasz, int ignore_index,
                                         float* out_scalars, int64_t* preds, int64_t* gt,
                                         float* dlogits, void* workspace, size_t workspace_bytes, void* stream);

/* Counters of tdssf_cels_fwd_bwd_counted since the library was loaded: graph replays, graph captures, plain launch sequences. */
TDSSF_API void tdssf_cels_graph_stats(int64_t* replays, int64_t* captures, int64_t* plain);

/* Backward through the individual terms (the reference's three loss outputs are all differentiable, utils/losses.py:60-70):
 * dlogits = coef[0] * d ce / d logits + coef[1] * d lovasz / d logits, from the workspace a tdssf_cels_fwd_bwd call on the
 * same n_rows left behind (probabilities, Jaccard gradients, per-class counts; the caller keeps it alive and unmodified).
 * coef: device float[2].  n_valid: the value that call returned in n_valid_host.  dlogits [n_rows,16] is overwritten. */
TDSSF_API int tdssf_cels_backward(const void* workspace, size_t workspace_bytes, int64_t n_rows, int64_t n_valid,
                                  const float* class_w, const float* coef, float* dlogits, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a11  confusion matrix
 * replaces  compute_confusion_matrix   utils/evaluation.py:35-42
 * preds, labels [n] int64 in 1..C; cm [C*C] int64, rows = truth; ACCUMULATES into cm. */
TDSSF_API int tdssf_confusion(const int64_t* preds, const int64_t* labels, int64_t n, int C, int64_t* cm,
                    void* stream);

/* ---------------------------------------------------------------------------------------------
 * f4  calibration composition and multi-sweep accumulation (input preparation upstream of a1)
 * replaces  the numpy block of CustomNuScenesDataset.__getitem__   utils/dataloader.py:123-152
 *           (lidar2cam = inv(T_cam->ego) @ T_lidar->ego in float64, cast to fp32 at utils/dataloader.py:226)
 * A pose is 7 doubles: quaternion (w,x,y,z) + translation (x,y,z) — a nuScenes calibrated_sensor / ego_pose record.
 * lidar_sensor [B,7], cam_sensor [B,NC,7].  lidar_ego [B,7] and cam_ego [B,NC,7] are optional (both or neither):
 * without them the result is the reference's two-stage product; with them the full chain
 * lidar -> ego(t_lidar) -> global -> ego(t_cam) -> camera (the records the reference fetches and ignores).
 * Rt_f32 [B,NC,16] row-major is the `Rt` input of tdssf_project / tdssf_fusion_fwd; Rt_f64 optional. */
TDSSF_API int tdssf_compose_lidar2cam(const double* lidar_sensor, const double* cam_sensor, const double* lidar_ego,
                                      const double* cam_ego, int B, int NC, float* Rt_f32, double* Rt_f64, void* stream);

/* Multi-sweep cloud (BASELINE configs[3]); the reference loads one sweep (utils/dataloader.py:63-77), the accumulation
 * follows nuscenes-devkit LidarPointCloud.from_file_multisweep (dependency pinned in the reference's environment.yml,
 * source not vendored): T[s] = inv(ref_sensor) . inv(ref_ego) . sweep_ego[s] . sweep_sensor[s], float64.
 * sweep_sensor, sweep_ego [S,7]; ref_sensor, ref_ego [7]; T_f64 [S,16].  S <= 64. */
TDSSF_API int tdssf_compose_sweep2ref(const double* sweep_sensor, const double* sweep_ego, const double* ref_sensor,
                                      const double* ref_ego, int S, double* T_f64, void* stream);
/* pts [S][P] rows of (x,y,z,intensity) fp32, `sweep_stride_floats` apart; counts [S] int32 (device) valid rows per sweep;
 * out: the transformed rows packed in sweep order (capacity S*P rows); out_total (device, optional) = sum of counts.
 * xyz are transformed in float64 and stored as fp32, intensity is copied. */
TDSSF_API int tdssf_accumulate_sweeps(const float* pts, int64_t sweep_stride_floats, const int32_t* counts, const double* T_f64,
                                      int S, int P, float* out, int32_t* out_total, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Multi-GPU: mean all-reduce of the fusion-MLP gradient bucket over NVLink peer memory - one kernel per step
 * in place of a library collective (the reference is single-process: no counterpart; semantics per SURVEY 8e: sum over
 * ranks / world size).  One process per GPU on ONE node; every rank creates an exchange buffer, the 64-byte handles are
 * exchanged by the host side (torch.distributed in tdssf/parallel.py) and opened by the peers.
 *   tdssf_p2p_create / _open / _close / _destroy   one-time set-up and tear-down (create allocates and synchronises)
 *   tdssf_p2p_allreduce_mean   grad[0..n_floats) := mean over ranks, in place, ONE launch on `stream`, no allocation, no host
 *                              synchronisation.  bufs_host: `world` pointers, this process's mappings of every rank's buffer
 *                              (its own at index `rank`).  epoch: the group's call count from 1, identical on all ranks.
 *                              n_floats % 4 == 0, n_floats <= max_floats, grad 16-byte aligned, world <= 16.
 *                              Every rank must call it the same number of times; the sum is formed in rank order on every
 *                              rank (bit-identical results across ranks). */
TDSSF_API size_t tdssf_p2p_buffer_bytes(int64_t max_floats);
TDSSF_API int tdssf_p2p_create(int64_t max_floats, void** buf_out, void* handle_out_host /* 64 bytes */);
TDSSF_API int tdssf_p2p_open(const void* handle_host /* 64 bytes */, void** peer_buf_out);
TDSSF_API int tdssf_p2p_close(void* peer_buf);
TDSSF_API int tdssf_p2p_destroy(void* own_buf);
TDSSF_API int tdssf_p2p_allreduce_mean(void* const* bufs_host, int world, int rank, float* grad, int64_t n_floats, int64_t max_floats,
                                       uint32_t epoch, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TDSSF_H_ */
