/*
 * ghost_b200 -- C ABI of the B200-native GHOST face-swap inference hot path.
 *
 * Every entry point takes plain pointers and sizes; device pointers are raw CUDA
 * device addresses (tensor.data_ptr()), `stream` is a cudaStream_t passed as void*.
 * All calls are asynchronous on `stream` unless stated otherwise and return
 * 0 on success or a negative ghost_status; ghost_last_error() gives the message
 * of the last failure on the calling thread.  There is no CPU fallback.
 *
 * Reference interfaces replaced (paths relative to the postworthy/ghost tree):
 *   network/AEI_Net.py:143-159               AEI_Net.__init__/forward/get_attr
 *   arcface_model/iresnet.py (un-vendored)   iresnet100(fp16=False).forward
 *   utils/inference/core.py:13-26            transform_target_to_torch
 *   utils/inference/faceshifter_run.py:5-23  faceshifter_batch
 *   utils/inference/image_processing.py:23-48,51-77   normalize_and_torch(_batch), get_final_image
 *   utils/inference/video_processing.py:191-243       get_final_video (compositing only)
 *   utils/inference/video_processing.py:246-285       Frames / face_enhancement
 *   utils/inference/masks.py:39-108                   face_mask_static / erode_and_blur
 *   utils/inference/video_processing.py:159-188       crop extraction (warpAffine + resize_frames)
 *   models/pix2pix_model.py:42-62, models/networks/generator.py:93-137,390-400   Pix2PixModel 'inference2' -> LIPSPADEGenerator
 */
#ifndef GHOST_B200_H_
#define GHOST_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GHOST_B200_VERSION 100

typedef enum {
  GHOST_OK = 0,
  GHOST_ERR_INVALID = -1, /* bad argument / unsupported configuration */
  GHOST_ERR_WEIGHTS = -2, /* missing or mis-shaped state_dict entry */
  GHOST_ERR_CUDA = -3,    /* CUDA runtime / driver error */
  GHOST_ERR_KERNEL = -4   /* device-side pipeline timeout (a kernel bug, never ignored) */
} ghost_status;

/* element types of buffers crossing the ABI */
typedef enum { GHOST_F32 = 0, GHOST_BF16 = 1, GHOST_F16 = 2, GHOST_U8 = 3 } ghost_dtype;

/* arithmetic mode of a network (= storage type of its activations and weights on the device):
 *   GHOST_F32  = CUDA-core fp32 correctness mode (1e-4 contract);
 *   GHOST_F16  = tcgen05 fp16 x fp16 -> fp32 production mode: the reference's own deployment dtype
 *                (inference.py:30 `G.half()`), meets the 1e-2 max-abs / 1e-3 mean-abs contract on Y;
 *   GHOST_BF16 = tcgen05 bf16 x bf16 -> fp32, same speed, wider range, 8 mantissa bits (outside that contract for
 *                random-init weights). */

typedef struct ghost_net ghost_net_t;

int ghost_version(void);
const char* ghost_last_error(void);
/* Returns GHOST_ERR_KERNEL if any kernel launched so far reported a device-side pipeline timeout
 * (meaningful after the caller synchronised the stream); every other entry point performs the same check. */
int ghost_check_errors(void);

/* ---- AEI_Net(backbone, num_blocks, c_id)  (network/AEI_Net.py:143-151).
 * backbone: "unet" or "linknet" ("resnet" -> GHOST_ERR_INVALID). */
int ghost_aei_create(ghost_net_t** out, const char* backbone, int num_blocks, int c_id, int mode);

/* ---- iresnet100(fp16=False) (inference.py:33-36) */
int ghost_arcface_create(ghost_net_t** out, int mode);

/* Hand over one state_dict entry under the reference's key name (fp32, host memory, contiguous).
 * The data is copied; unknown keys are rejected at finalize time only if a required key is missing. */
int ghost_net_set_weight(ghost_net_t* net, const char* ref_key, const float* host_data, const int64_t* shape,
                         int ndim);

/* Fold BatchNorm, repack into kernel layouts, upload.  Synchronous. */
int ghost_net_finalize(ghost_net_t* net);

/* Y, attr = G(Xt, z_id)  (network/AEI_Net.py:153-156).
 *   Xt   [B,3,256,256] NCHW contiguous, xt_dtype in {F32,BF16,F16}, RGB in [-1,1]
 *   z_id [B,c_id] (or [B,1,c_id]) same set of dtypes
 *   Y    [B,3,256,256] NCHW out_dtype in {F32,BF16,F16}
 *   attr_out: NULL, or 8 device pointers (each may be NULL) receiving z_attr1..8 in NCHW out_dtype */
int ghost_aei_forward(ghost_net_t* net, const void* Xt, int xt_dtype, const void* z_id, int z_dtype, void* Y,
                      void* const* attr_out, int out_dtype, int B, void* stream);

/* attr = G.get_attr(X)  (network/AEI_Net.py:158-159) */
int ghost_aei_encode(ghost_net_t* net, const void* X, int x_dtype, void* const* attr_out, int out_dtype, int B,
                     void* stream);

/* Fused inference entry (core.py:13-26 + faceshifter_run.py:5-23 with the bounce through the host removed):
 *   crops_bgr [B,256,256,3] uint8 BGR (device) -> normalise -> G -> optional uint8 BGR [B,256,256,3] (device).
 *   z_id: [1,c_id] (broadcast, as faceshifter_batch tiles source_emb) when z_rows == 1, else [B,c_id].
 *   out_bgr may be NULL when only the on-device Y (ghost_net_y_nhwc) is wanted by the paste-back. */
int ghost_aei_swap_u8(ghost_net_t* net, const uint8_t* crops_bgr, const void* z_id, int z_dtype, int z_rows,
                      uint8_t* out_bgr, int B, void* stream);

/* ghost_aei_swap_u8 in two halves, for callers that overlap the attribute encoder with the ArcFace network (the
 * encoder, network/AEI_Net.py:72-95, does not depend on z_id): encode_u8 = normalise + MLAttrEncoder on `stream`;
 * generate_u8 = AADGenerator (+ optional uint8 conversion) on `stream`, which must be ordered after encode_u8 of
 * the same batch size (an event between the two streams).  Results are identical to ghost_aei_swap_u8. */
int ghost_aei_encode_u8(ghost_net_t* net, const uint8_t* crops_bgr, int B, void* stream);
int ghost_aei_generate_u8(ghost_net_t* net, const void* z_id, int z_dtype, int z_rows, uint8_t* out_bgr, int B,
                          void* stream);

/* Device pointer + dtype of the NHWC-3 generator output of the last forward with batch B (valid until the next call). */
int ghost_net_y_nhwc(ghost_net_t* net, int B, void** ptr, int* dtype);

/* z = netArc(x), x [B,3,112,112] NCHW -> z [B,512], not normalised (core.py:45,54). */
int ghost_arcface_forward(ghost_net_t* net, const void* x, int x_dtype, void* z, int z_dtype, int B, void* stream);

/* normalize_and_torch(_batch) + F.interpolate(0.5, bilinear, align_corners=True) + netArc in one call:
 * src [B,224,224,3] RGB HWC, uint8 (is_u8=1, divided by 255) or float32 in [0,1] (device). */
int ghost_arcface_embed_hwc(ghost_net_t* net, const void* src, int is_u8, void* z, int z_dtype, int B, void* stream);

/* Capture the forward of batch size B into a CUDA graph (optional; replayed by later calls). */
int ghost_net_enable_graph(ghost_net_t* net, int enable);

/* number of kernel launches one forward of batch B issues, and its dense FLOPs (for bench accounting) */
int ghost_net_stats(ghost_net_t* net, int B, int* launches, double* flops);

/* Per-op timing of one forward at batch B (the network must have run at batch B at least once, so that its input
 * buffers hold data): every op is launched `reps` times bracketed by CUDA events on `stream`.  Synchronous.
 * For each op i < *n_ops (at most max_ops): ms[i] mean duration, flops[i] dense 2*MAC flops, bytes[i] algorithmic
 * HBM bytes (bandwidth-bound ops), cls[i] kernel class (0 tcgen05 conv, 1 tcgen05 gamma/beta GEMM + fused AAD blend,
 * 2 CUDA-core conv, 3 instance-norm statistics, 4 bilinear x2, 5 misc), names + i*name_stride a NUL-terminated name. */
int ghost_net_profile(ghost_net_t* net, int B, int reps, int max_ops, float* ms, double* flops, double* bytes, int* cls,
                      char* names, int name_stride, int* n_ops, void* stream);

/* Block-level parity hook: copy a named intermediate activation of the last forward at batch B ("m1".."m8" block
 * inputs, "y1".."y8" block outputs, "a1".."a8" last AAD+ReLU output of a block) into dst as NCHW dst_dtype.
 * shape4 (optional) receives {N,C,H,W}; dst may be NULL to query the shape only. */
int ghost_net_debug_get(ghost_net_t* net, int B, const char* name, void* dst, int dst_dtype, int* shape4,
                        void* stream);

/* ---- pix2pix super-resolution generator: LIPSPADEGenerator(TestOptions) as inference.py:42-48 builds it
 * (models/pix2pix_model.py:16-25 -> models/networks/generator.py:390-400, config models/config_sr.py:3-29), in the
 * `.train()` state of inference.py:48: every SPADE normalises with the statistics of the current batch and every
 * spectral-normalised conv performs one power iteration per forward (its weight_u / weight_v buffers advance).
 * Weights go in through ghost_net_set_weight under the reference's state_dict keys (head_0.conv_0.weight_orig,
 * head_0.conv_0.weight_u, ups.3.norm_s.mlp_gamma.weight, lip_encoder.model.3.logit.1.bias, ...). */
int ghost_sr_create(ghost_net_t** out, int mode);

/* generated = netG(x)  (Pix2PixModel.forward(data, 'inference2'), models/pix2pix_model.py:58-61;
 * SPADEGenerator.forward, models/networks/generator.py:93-137 with seg = input).
 *   x [B,3,256,256] NCHW RGB in [0,1], y [B,3,256,256] NCHW in [-1,1] (tanh).  The B crops share their batch statistics. */
int ghost_sr_forward(ghost_net_t* net, const void* x, int x_dtype, void* y, int y_dtype, int B, void* stream);

/* One DataLoader batch of face_enhancement (utils/inference/video_processing.py:263-285, batch_size = 20):
 *   crops_bgr [B,256,256,3] uint8 BGR (device) -> RGB / 255 -> netG -> clamp(g*255, 0, 255) -> uint8 -> BGR. */
int ghost_sr_enhance_u8(ghost_net_t* net, const uint8_t* crops_bgr, uint8_t* out_bgr, int B, void* stream);

/* Reads a spectral-norm buffer of the finalized network ("head_0.conv_0.weight_u", "...weight_v") into host memory
 * (what state_dict() of the reference module would hold after the forwards run so far).  Synchronous.
 * dst_host may be NULL to query the length (*n_out). */
int ghost_sr_read_buffer(ghost_net_t* net, const char* key, float* dst_host, int cap, int* n_out, void* stream);

void ghost_net_destroy(ghost_net_t* net);

/* ---- standalone conversions (device buffers) */
int ghost_preproc_u8(const uint8_t* crops_bgr, void* dst_nhwc3, int dst_dtype, int B, int half_rounding, void* stream);
int ghost_postproc_u8(const void* y_nhwc3, int y_dtype, uint8_t* dst_bgr, int B, void* stream);

/* ---- paste-back (image_processing.py:51-77 exact mode; video_processing.py:212-233 float mode) */
typedef enum { GHOST_PASTE_EXACT = 0, GHOST_PASTE_FLOAT = 1 } ghost_paste_mode;
typedef enum { GHOST_CROP_U8_BGR = 0, GHOST_CROP_Y_BF16 = 1, GHOST_CROP_Y_F32 = 2, GHOST_CROP_Y_F16 = 3 } ghost_crop_kind;

typedef struct {
  const void* crops;        /* device: [n_crops,256,256,3] u8 BGR, or NHWC-3 generator output (RGB, [-1,1]) */
  int crop_kind;            /* ghost_crop_kind */
  int n_crops;
  const void* masks;        /* device: [n_masks,224,224] float64 (mask_f64=1) or float32 */
  int mask_f64;
  int n_masks;
  const uint8_t* frames_in; /* device: [n_frames,H,W,3] u8 BGR */
  uint8_t* frames_out;      /* device: [n_frames,H,W,3]; may alias frames_in */
  int n_frames, H, W;
  int n_faces;
  const double* tfm;        /* host: [n_faces][2][3] frame->224-crop affines (tfm_arrays) */
  const int* face_frame;    /* host: [n_faces] frame index of each face, non-decreasing; faces of one frame
                               compose in list order */
  const int* face_crop;     /* host: [n_faces] crop index */
  const int* face_mask;     /* host: [n_faces] mask index */
  int mode;                 /* ghost_paste_mode */
} ghost_paste_desc;

/* Re-entrant: the per-call tables and the resized crops live in a stream-ordered allocation (cudaMallocAsync on
 * `stream`, released behind the kernels), so concurrent calls on different streams / threads share nothing. */
int ghost_paste_back(const ghost_paste_desc* desc, void* stream);

/* The paste-back rewrites only the bounding box of each face (the image of the 224 x 224 crop under the inverse of
 * its affine, clipped to the frame; image_processing.py:66-75 composites `mask*swap + (1-mask)*frame`, which is the
 * frame itself wherever the warped mask is 0).  ghost_paste_boxes returns those boxes, boxes[i] = {x0, y0, x1, y1}
 * (x1, y1 exclusive; all 0 for a face outside its frame), so that a caller who keeps the frames in pinned host
 * memory can move just the boxes: ghost_copy_boxes copies the box pixels of [n_frames,H,W,3] u8 frames between a
 * host and a device buffer of the same layout (to_device != 0: host -> device), asynchronously on `stream`. */
int ghost_paste_boxes(const double* tfm, int n_faces, int H, int W, int* boxes);
int ghost_copy_boxes(void* dst, const void* src, int n_frames, int H, int W, const int* face_frame, const int* boxes,
                     int n_faces, int to_device, void* stream);

/* ---- identity matching for multi-face video (utils/inference/video_processing.py:126-148):
 *   similarity = F.normalize(face_embeds) @ F.normalize(target_embeds).T;  best_idx = similarity.argmax(0);
 *   matched[j] = similarity[best_idx[j]][j] > similarity_th.
 * face_embeds [n_faces][dim], target_embeds [n_targets][dim] fp32 un-normalised ArcFace outputs (device);
 * best_idx / best_sim / matched [n_targets] (device). */
int ghost_match_identities(const float* face_embeds, int n_faces, const float* target_embeds, int n_targets, int dim,
                           float similarity_th, int* best_idx, float* best_sim, int* matched, void* stream);

/* ---- crop extraction in front of the generator (utils/inference/video_processing.py:159-168 + resize_frames
 * :174-188; image_processing.py:13-20), bit-exact against the reference's OpenCV calls:
 *   crops224[i] = cv2.warpAffine(frames[face_frame[i]], tfm[i], (224, 224), borderValue=0.0)
 *   crops256[i] = cv2.resize(crops224[i], (256, 256))
 * frames [n_frames,H,W,3] u8 BGR (device); tfm [n][2][3] float64 frame->crop affines (host, tfm_array entries);
 * face_frame [n] (host); crops224 [n,224,224,3] / crops256 [n,256,256,3] u8 on the device, either may be NULL. */
int ghost_crop_faces(const uint8_t* frames, int n_frames, int H, int W, const double* tfm, const int* face_frame, int n,
                     uint8_t* crops224, uint8_t* crops256, void* stream);

/* ---- ghost_crop_faces without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same kernels and bit-identical results as ghost_crop_faces, but the scratch (face
 * table, resize tables, the 224x224 intermediate when crops224 is NULL) lives in one file-scope
 * `static __device__` array of GHOST_CROP_NOALLOC_MAX_FACES * 150528 + 6144 bytes (~4.6 MiB, held by every GPU that
 * loads the library until exit) whose device address is resolved once per device with cudaGetSymbolAddress.  The call
 * performs no cudaMalloc / cudaMallocAsync / cudaFree and no allocating library routine; the caller passes every input
 * and output buffer.  n > GHOST_CROP_NOALLOC_MAX_FACES -> GHOST_ERR_INVALID with a message naming the limit.
 * All calls share the array: calls that overlap on the GPU (different streams or threads) race - issue them on one
 * stream or order them with events.  Nothing in the array needs to start at zero.  The first call on a device does
 * the set-up (one synchronous 3.5 KB copy): make it before a graph capture begins. */
#define GHOST_CROP_NOALLOC_MAX_FACES 32
int ghost_crop_faces_noalloc(const uint8_t* frames, int n_frames, int H, int W, const double* tfm, const int* face_frame,
                             int n, uint8_t* crops224, uint8_t* crops256, void* stream);

/* ---- peer-mapped gather buffers: the one exchange of the sharded path, the final gather of finished
 * uint8 frames on the consumer rank, as copy-engine transfers into a buffer the consumer exports through CUDA IPC.
 * No counterpart in the reference (it has no multi-GPU inference path); ghost_b200/shard.py::PeerGather is the host
 * side.  ghost_peer_alloc: cudaMalloc (zeroed) + the 64-byte IPC handle of the allocation (consumer rank);
 * ghost_peer_open / ghost_peer_close: map / unmap it in another process of the same box; ghost_peer_copy: asynchronous
 * copy (cudaMemcpyDefault) on `stream`; ghost_peer_signal: stores `value` in a 32-bit flag (possibly peer memory)
 * behind the work queued on `stream`; ghost_peer_wait: makes `stream` wait until *flag >= value (cuStreamWaitValue32,
 * no kernel). */
int ghost_peer_alloc(size_t bytes, void** dev_ptr, unsigned char* handle64);
int ghost_peer_open(const unsigned char* handle64, void** dev_ptr);
int ghost_peer_close(void* dev_ptr);
int ghost_peer_free(void* dev_ptr);
int ghost_peer_copy(void* dst, const void* src, size_t bytes, void* stream);
int ghost_peer_signal(uint32_t* flag, uint32_t value, void* stream);
int ghost_peer_wait(uint32_t* flag, uint32_t value, void* stream);

/* ---- blend-mask synthesis (utils/inference/masks.py:39-108 face_mask_static + erode_and_blur, called at
 * image_processing.py:68 and video_processing.py:221-223), bit-exact against the reference's OpenCV calls:
 *   landmarks [n][106][2] float32 (host): the 106-point landmarks of the swapped 224x224 crop, as
 *             Handler.get_without_detection_without_transform returns them;
 *   params    [n][3] int (host): erode, sigmaX, sigmaY - the preset face_mask_static picked (masks.py:43-69) or the
 *             caller's `params`; erode < 0 dilates;
 *   masks_out [n][224][224] on the device, float64 (mask_f64 = 1, get_final_image) or float32 (get_final_video),
 *             the layout ghost_paste_desc.masks reads.
 * Eyebrow lift, convex hull and polygon rasterisation run on the host (integer work on 106 points), erode / dilate,
 * border clip and the fixed-point Gaussian blur on the device, one CTA per face. */
int ghost_face_masks(const float* landmarks, const int* params, int n, void* masks_out, int mask_f64, void* stream);

/* ---- single-op entry points used by the kernel-level parity tests (tests/test_kernels_gpu.py).
 * conv: NHWC activations (dtype = mode); weights fp32 on the host in the reference layout:
 * kind 0 = nn.Conv2d(k,stride,pad) weight [Cout][Cin][k][k], kind 1 = nn.ConvTranspose2d(k=4,s=2,p=1) weight
 * [Cin][Cout][4][4].  act: 0 none, 1 leaky(slope), 3 relu.  mode GHOST_F32 / GHOST_BF16 / GHOST_F16.  Synchronous. */
int ghost_test_conv(const void* x_nhwc, int N, int H, int W, int Cin, const float* w_host, const float* bias_host,
                    int Cout, int kind, int k, int stride, int pad, int act, float slope, void* y_nhwc, int mode,
                    int bn_hint, void* stream);
/* AADLayer + ReLU on NHWC tensors (network/AADLayer.py:20-38): weights on the host in reference layout. */
int ghost_test_aad(const void* h_nhwc, const void* zattr_nhwc, const float* zid_host, int N, int H, int W, int c_x,
                   int c_attr, int c_id, const float* conv1_w, const float* conv1_b, const float* conv2_w,
                   const float* conv2_b, const float* fc1_w, const float* fc1_b, const float* fc2_w,
                   const float* fc2_b, const float* convh_w, float convh_b, void* out_nhwc, int mode, void* stream);
const char* ghost_test_last_error(void);
/* mean duration (ms) of the re-launches ghost_test_conv timed when GHOST_TEST_CONV_REPS=n is set (tools/conv_sweep.py) */
float ghost_test_last_ms(void);
int ghost_test_upsample2x(const void* x_nhwc, void* y_nhwc, int dtype, int N, int H, int W, int C, void* stream);
/* Bilinear x2 fused with the InstanceNorm statistics of its output (the next AAD block's input,
 * network/AEI_Net.py:124-138 + network/AADLayer.py:21); mean / biased variance per (n, c) come back on the host. */
int ghost_test_upsample2x_stats(const void* x_nhwc, void* y_nhwc, int dtype, int N, int H, int W, int C,
                                float* mean_host, float* var_host, void* stream);

/* Host half of ghost_face_masks for one face (no device work): hull bitmap (224 rows x 7 words, bit x of row y =
 * pixel (y, x)) and the fixed-point Gaussian kernels (128 entries each, nxny = {nx, ny} taps used). */
int ghost_test_mask_job(const float* landmarks, int erode, int sigma_x, int sigma_y, uint32_t* bitmap, uint16_t* kx,
                        uint16_t* ky, int* nxny);

#ifdef __cplusplus
}
#endif
#endif /* GHOST_B200_H_ */
