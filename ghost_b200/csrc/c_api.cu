// extern "C" surface declared in include/ghost_b200.h.  No C++ exception crosses the ABI.
#include <string.h>

#include <mutex>

#include "../../include/ghost_b200.h"
#include "net.cuh"
#include "sr_kernels.cuh"
#include "mask.cuh"
#include "crop.cuh"
#include "match.cuh"

using namespace ghost;

namespace {
thread_local std::string g_last_error;
}
namespace ghost {
void set_last_error(const char* msg) { g_last_error = msg ? msg : ""; }
}  // namespace ghost

struct ghost_net {
  Net net;
  std::mutex mu;
  ghost_net(Net::Kind k, const char* bb, int nb, int cid, int mode) : net(k, bb, nb, cid, mode) {}
};

namespace {

template <typename F>
int guarded(F&& f) {
  try {
    check_device_errors();
    f();
    return GHOST_OK;
  } catch (const GhostError& e) {
    g_last_error = e.what();
    return e.code;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return GHOST_ERR_INVALID;
  } catch (...) {
    g_last_error = "unknown error";
    return GHOST_ERR_INVALID;
  }
}

inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

void require(bool c, const char* msg) {
  if (!c) fail(GHOST_ERR_INVALID, msg);
}

void write_attrs(Net& n, ExecPlan& p, void* const* attr_out, int out_dtype, cudaStream_t s) {
  if (!attr_out) return;
  for (int k = 0; k < 8; ++k) {
    if (!attr_out[k]) continue;
    const Act& a = p.attr[k];
    check_msg(nhwc_to_nchw_launch(a.p, a.dtype, a.C, 0, attr_out[k], out_dtype, a.N, a.C, a.H, a.W, s), "attr out");
  }
}

struct DeviceTables {
  short* idx = nullptr;
  short* a0 = nullptr;
  short* a1 = nullptr;
};
DeviceTables& resize_tables() {
  static std::map<int, DeviceTables> per_dev;
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  int dev = 0;
  check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  DeviceTables& t = per_dev[dev];
  if (!t.idx) {
    short h[3][224];
    paste_resize_tables(h[0], h[1], h[2]);
    short* d = nullptr;
    check_cuda(cudaMalloc(&d, sizeof(h)), "resize tables");
    check_cuda(cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice), "resize tables");
    t.idx = d;
    t.a0 = d + 224;
    t.a1 = d + 448;
  }
  return t;
}

// Per-call scratch (face tables, resized crops) comes from the device's stream-ordered pool, so that calls on
// different streams or threads never share a buffer.  The pool keeps what it has handed out once
// (release threshold = max): after the first call an allocation is a pointer bump on the stream, not a cudaMalloc.
void* stream_scratch(size_t bytes, cudaStream_t s) {
  static std::mutex mu;
  static bool tuned[64] = {};
  int dev = 0;
  check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  {
    std::lock_guard<std::mutex> lk(mu);
    if (dev >= 0 && dev < 64 && !tuned[dev]) {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
      }
      cudaGetLastError();
      tuned[dev] = true;
    }
  }
  void* p = nullptr;
  check_cuda(cudaMallocAsync(&p, bytes, s), "stream-ordered scratch");
  return p;
}

}  // namespace

extern "C" {

int ghost_version(void) { return GHOST_B200_VERSION; }
const char* ghost_last_error(void) { return g_last_error.c_str(); }

int ghost_check_errors(void) {
  return guarded([] {});
}

int ghost_aei_create(ghost_net_t** out, const char* backbone, int num_blocks, int c_id, int mode) {
  return guarded([&] {
    require(out && backbone, "null argument");
    require(mode == GHOST_F32 || mode == GHOST_BF16 || mode == GHOST_F16, "mode must be GHOST_F32, GHOST_BF16 or GHOST_F16");
    if (strcmp(backbone, "resnet") == 0) fail(GHOST_ERR_INVALID, "backbone 'resnet' is not supported by ghost_b200");
    require(strcmp(backbone, "unet") == 0 || strcmp(backbone, "linknet") == 0, "unknown backbone");
    require(num_blocks >= 1 && num_blocks <= 4, "num_blocks must be in [1,4]");
    require(c_id > 0 && c_id % 32 == 0, "c_id must be a positive multiple of 32");
    *out = new ghost_net(Net::kAEI, backbone, num_blocks, c_id, mode);
  });
}

int ghost_arcface_create(ghost_net_t** out, int mode) {
  return guarded([&] {
    require(out, "null argument");
    require(mode == GHOST_F32 || mode == GHOST_BF16 || mode == GHOST_F16, "mode must be GHOST_F32, GHOST_BF16 or GHOST_F16");
    *out = new ghost_net(Net::kArcFace, "iresnet100", 0, 512, mode);
  });
}

int ghost_net_set_weight(ghost_net_t* net, const char* key, const float* data, const int64_t* shape, int ndim) {
  return guarded([&] {
    require(net && key && data && (shape || ndim == 0) && ndim >= 0 && ndim <= 8, "bad argument");
    std::lock_guard<std::mutex> lk(net->mu);
    net->net.set_weight(key, data, shape, ndim);
  });
}

int ghost_net_finalize(ghost_net_t* net) {
  return guarded([&] {
    require(net, "null net");
    std::lock_guard<std::mutex> lk(net->mu);
    net->net.finalize();
  });
}

int ghost_net_enable_graph(ghost_net_t* net, int enable) {
  return guarded([&] {
    require(net, "null net");
    net->net.use_graph = enable != 0;
  });
}

int ghost_net_stats(ghost_net_t* net, int B, int* launches, double* flops) {
  return guarded([&] {
    require(net, "null net");
    std::lock_guard<std::mutex> lk(net->mu);
    ExecPlan& p = net->net.plan(B);
    if (launches) *launches = static_cast<int>(p.ops.size());
    if (flops) *flops = p.flops;
  });
}

int ghost_aei_forward(ghost_net_t* net, const void* Xt, int xt_dtype, const void* z_id, int z_dtype, void* Y,
                      void* const* attr_out, int out_dtype, int B, void* stream) {
  return guarded([&] {
    require(net && Xt && z_id && Y, "null argument");
    require(net->net.kind == Net::kAEI, "not an AEI_Net handle");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    check_msg(nchw_to_nhwc_launch(Xt, xt_dtype, p.x_in.p, p.x_in.dtype, B, 3, 256, 256, 3, 0, s), "Xt in");
    check_msg(nchw_to_nhwc_launch(z_id, z_dtype, p.zid.p, p.zid.dtype, B, n.c_id, 1, 1, n.c_id, 0, s), "z_id in");
    n.run_plan(p, s);
    check_msg(nhwc_to_nchw_launch(p.y.p, p.y.dtype, 3, 0, Y, out_dtype, B, 3, 256, 256, s), "Y out");
    write_attrs(n, p, attr_out, out_dtype, s);
  });
}

int ghost_aei_encode(ghost_net_t* net, const void* X, int x_dtype, void* const* attr_out, int out_dtype, int B,
                     void* stream) {
  return guarded([&] {
    require(net && X && attr_out, "null argument");
    require(net->net.kind == Net::kAEI, "not an AEI_Net handle");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    check_msg(nchw_to_nhwc_launch(X, x_dtype, p.x_in.p, p.x_in.dtype, B, 3, 256, 256, 3, 0, s), "X in");
    n.run_plan(p, s, 0, n.encoder_ops);
    write_attrs(n, p, attr_out, out_dtype, s);
  });
}

int ghost_aei_swap_u8(ghost_net_t* net, const uint8_t* crops_bgr, const void* z_id, int z_dtype, int z_rows,
                      uint8_t* out_bgr, int B, void* stream) {
  return guarded([&] {
    require(net && crops_bgr && z_id, "null argument");
    require(net->net.kind == Net::kAEI, "not an AEI_Net handle");
    require(z_rows == 1 || z_rows == B, "z_id must have 1 or B rows");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    const long long npix = static_cast<long long>(B) * 256 * 256;
    // bf16 mode rounds once from fp32; the fp32 mode keeps full precision (core.py:13-26)
    check_msg(preproc_u8_launch(crops_bgr, p.x_in.p, p.x_in.dtype, npix, 0, s), "preproc");
    const size_t zb = (p.zid.dtype == kF32 ? 4 : 2) * static_cast<size_t>(n.c_id);
    if (z_rows == B) {
      check_msg(nchw_to_nhwc_launch(z_id, z_dtype, p.zid.p, p.zid.dtype, B, n.c_id, 1, 1, n.c_id, 0, s), "z_id in");
    } else {
      // torch.cat([source_emb]*bs) (faceshifter_run.py:15-16)
      check_msg(nchw_to_nhwc_launch(z_id, z_dtype, p.zid.p, p.zid.dtype, 1, n.c_id, 1, 1, n.c_id, 0, s), "z_id in");
      for (int filled = 1; filled < B;) {
        const int cnt = filled < B - filled ? filled : B - filled;
        check_cuda(cudaMemcpyAsync(static_cast<char*>(p.zid.p) + zb * filled, p.zid.p, zb * cnt,
                                   cudaMemcpyDeviceToDevice, s),
                   "z_id tile");
        filled += cnt;
      }
    }
    n.run_plan(p, s);
    if (out_bgr) check_msg(postproc_u8_launch(p.y.p, p.y.dtype, out_bgr, npix, s), "postproc");
  });
}

int ghost_aei_encode_u8(ghost_net_t* net, const uint8_t* crops_bgr, int B, void* stream) {
  return guarded([&] {
    require(net && crops_bgr, "null argument");
    require(net->net.kind == Net::kAEI, "not an AEI_Net handle");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    check_msg(preproc_u8_launch(crops_bgr, p.x_in.p, p.x_in.dtype, static_cast<long long>(B) * 65536, 0, s), "preproc");
    n.run_plan(p, s, 0, n.encoder_ops);
  });
}

int ghost_aei_generate_u8(ghost_net_t* net, const void* z_id, int z_dtype, int z_rows, uint8_t* out_bgr, int B,
                          void* stream) {
  return guarded([&] {
    require(net && z_id, "null argument");
    require(net->net.kind == Net::kAEI, "not an AEI_Net handle");
    require(z_rows == 1 || z_rows == B, "z_id must have 1 or B rows");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    const size_t zb = (p.zid.dtype == kF32 ? 4 : 2) * static_cast<size_t>(n.c_id);
    check_msg(nchw_to_nhwc_launch(z_id, z_dtype, p.zid.p, p.zid.dtype, z_rows == B ? B : 1, n.c_id, 1, 1, n.c_id, 0, s),
              "z_id in");
    for (int filled = 1; z_rows == 1 && filled < B;) {
      const int cnt = filled < B - filled ? filled : B - filled;
      check_cuda(cudaMemcpyAsync(static_cast<char*>(p.zid.p) + zb * filled, p.zid.p, zb * cnt, cudaMemcpyDeviceToDevice, s),
                 "z_id tile");
      filled += cnt;
    }
    n.run_plan(p, s, n.encoder_ops);
    if (out_bgr) check_msg(postproc_u8_launch(p.y.p, p.y.dtype, out_bgr, static_cast<long long>(B) * 65536, s), "postproc");
  });
}

int ghost_net_y_nhwc(ghost_net_t* net, int B, void** ptr, int* dtype) {
  return guarded([&] {
    require(net && ptr && dtype, "null argument");
    std::lock_guard<std::mutex> lk(net->mu);
    ExecPlan& p = net->net.plan(B);
    *ptr = p.y.p;
    *dtype = p.y.dtype;
  });
}

int ghost_arcface_forward(ghost_net_t* net, const void* x, int x_dtype, void* z, int z_dtype, int B, void* stream) {
  return guarded([&] {
    require(net && x && z, "null argument");
    require(net->net.kind == Net::kArcFace, "not an iresnet handle");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    check_msg(nchw_to_nhwc_launch(x, x_dtype, p.emb_in.p, p.emb_in.dtype, B, 3, 112, 112, 3, 0, s), "x in");
    n.run_plan(p, s);
    check_msg(nhwc_to_nchw_launch(p.emb_out, kF32, 512, 0, z, z_dtype, B, 512, 1, 1, s), "z out");
  });
}

int ghost_arcface_embed_hwc(ghost_net_t* net, const void* src, int is_u8, void* z, int z_dtype, int B, void* stream) {
  return guarded([&] {
    require(net && src && z, "null argument");
    require(net->net.kind == Net::kArcFace, "not an iresnet handle");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    check_msg(arcface_preproc_launch(src, is_u8, p.emb_in.p, p.emb_in.dtype, B, s), "arcface preproc");
    n.run_plan(p, s);
    check_msg(nhwc_to_nchw_launch(p.emb_out, kF32, 512, 0, z, z_dtype, B, 512, 1, 1, s), "z out");
  });
}

int ghost_net_profile(ghost_net_t* net, int B, int reps, int max_ops, float* ms, double* flops, double* bytes, int* cls,
                      char* names, int name_stride, int* n_ops, void* stream) {
  return guarded([&] {
    require(net && n_ops && reps > 0 && max_ops >= 0, "bad argument");
    std::lock_guard<std::mutex> lk(net->mu);
    ExecPlan& p = net->net.plan(B);
    std::vector<float> t;
    net->net.profile_plan(p, S(stream), reps, &t);
    const int n = static_cast<int>(p.ops.size());
    *n_ops = n;
    for (int i = 0; i < n && i < max_ops; ++i) {
      if (ms) ms[i] = t[i];
      if (flops) flops[i] = p.ops[i].flops;
      if (bytes) bytes[i] = p.ops[i].bytes;
      if (cls) cls[i] = p.ops[i].cls;
      if (names && name_stride > 0) {
        strncpy(names + static_cast<size_t>(i) * name_stride, p.ops[i].name.c_str(), name_stride - 1);
        names[static_cast<size_t>(i) * name_stride + name_stride - 1] = 0;
      }
    }
  });
}

int ghost_net_debug_get(ghost_net_t* net, int B, const char* name, void* dst, int dst_dtype, int* shape4,
                        void* stream) {
  return guarded([&] {
    require(net && name, "null argument");
    std::lock_guard<std::mutex> lk(net->mu);
    ExecPlan& p = net->net.plan(B);
    auto it = p.named.find(name);
    if (it == p.named.end()) fail(GHOST_ERR_INVALID, std::string("no intermediate named ") + name);
    const Act& a = it->second;
    if (shape4) {
      shape4[0] = a.N; shape4[1] = a.C; shape4[2] = a.H; shape4[3] = a.W;
    }
    if (dst)
      check_msg(nhwc_to_nchw_launch(a.p, a.dtype, a.C, 0, dst, dst_dtype, a.N, a.C, a.H, a.W, S(stream)), "debug get");
  });
}

int ghost_sr_create(ghost_net_t** out, int mode) {
  return guarded([&] {
    require(out, "null argument");
    require(mode == GHOST_F32 || mode == GHOST_BF16 || mode == GHOST_F16, "mode must be GHOST_F32, GHOST_BF16 or GHOST_F16");
    *out = new ghost_net(Net::kSR, "lipspade", 0, 0, mode);
  });
}

int ghost_sr_forward(ghost_net_t* net, const void* x, int x_dtype, void* y, int y_dtype, int B, void* stream) {
  return guarded([&] {
    require(net && x && y, "null argument");
    require(net->net.kind == Net::kSR, "not a super-resolution generator handle");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    check_msg(nchw_to_nhwc_launch(x, x_dtype, p.x_in.p, p.x_in.dtype, B, 3, 256, 256, 3, 0, s), "x in");
    n.run_plan(p, s);
    check_msg(nhwc_to_nchw_launch(p.y.p, p.y.dtype, 3, 0, y, y_dtype, B, 3, 256, 256, s), "y out");
  });
}

int ghost_sr_enhance_u8(ghost_net_t* net, const uint8_t* crops_bgr, uint8_t* out_bgr, int B, void* stream) {
  return guarded([&] {
    require(net && crops_bgr && out_bgr, "null argument");
    require(net->net.kind == Net::kSR, "not a super-resolution generator handle");
    std::lock_guard<std::mutex> lk(net->mu);
    Net& n = net->net;
    ExecPlan& p = n.plan(B);
    cudaStream_t s = S(stream);
    check_msg(sr_preproc_launch(crops_bgr, p.x_in.p, p.x_in.dtype, B, 65536, s), "sr preproc");
    n.run_plan(p, s);
    check_msg(sr_postproc_launch(p.y.p, p.y.dtype, out_bgr, B, 65536, s), "sr postproc");
  });
}

int ghost_sr_read_buffer(ghost_net_t* net, const char* key, float* dst_host, int cap, int* n_out, void* stream) {
  return guarded([&] {
    require(net && key, "null argument");
    require(net->net.kind == Net::kSR, "not a super-resolution generator handle");
    std::lock_guard<std::mutex> lk(net->mu);
    if (!net->net.finalized) fail(GHOST_ERR_INVALID, "network used before ghost_net_finalize");
    const int n = net->net.read_sn_buffer(key, dst_host, cap, S(stream));
    if (n_out) *n_out = n;
  });
}

void ghost_net_destroy(ghost_net_t* net) { delete net; }

int ghost_preproc_u8(const uint8_t* crops_bgr, void* dst, int dst_dtype, int B, int half_rounding, void* stream) {
  return guarded([&] {
    require(crops_bgr && dst && B > 0, "bad argument");
    check_msg(preproc_u8_launch(crops_bgr, dst, dst_dtype, static_cast<long long>(B) * 65536, half_rounding, S(stream)),
              "preproc");
  });
}

int ghost_postproc_u8(const void* y, int y_dtype, uint8_t* dst, int B, void* stream) {
  return guarded([&] {
    require(y && dst && B > 0, "bad argument");
    check_msg(postproc_u8_launch(y, y_dtype, dst, static_cast<long long>(B) * 65536, S(stream)), "postproc");
  });
}

int ghost_match_identities(const float* face_embeds, int n_faces, const float* target_embeds, int n_targets, int dim,
                           float similarity_th, int* best_idx, float* best_sim, int* matched, void* stream) {
  return guarded([&] {
    require(face_embeds && target_embeds && best_idx && best_sim && matched, "null argument");
    require(n_faces >= 0 && n_targets > 0 && dim > 0, "bad sizes");
    check_msg(match_identities_launch(face_embeds, n_faces, target_embeds, n_targets, dim, similarity_th, best_idx,
                                      best_sim, matched, S(stream)), "match identities");
  });
}

int ghost_crop_faces(const uint8_t* frames, int n_frames, int H, int W, const double* tfm, const int* face_frame, int n,
                     uint8_t* crops224, uint8_t* crops256, void* stream) {
  return guarded([&] {
    require(frames && tfm && face_frame && n > 0 && n_frames > 0 && H > 0 && W > 0, "bad argument");
    require(crops224 || crops256, "no output buffer");
    cudaStream_t s = S(stream);
    std::vector<CropFace> faces(n);
    for (int i = 0; i < n; ++i) {
      require(face_frame[i] >= 0 && face_frame[i] < n_frames, "face_frame out of range");
      crop_prepare_face(&faces[i], tfm + 6 * i, face_frame[i]);
    }
    short tables[7 * 256];
    crop_resize_tables(tables);
    const size_t fb = (faces.size() * sizeof(CropFace) + 255) / 256 * 256, tb = sizeof(tables);
    const size_t scratch224 = crops224 ? 0 : static_cast<size_t>(n) * 224 * 224 * 3;
    char* dev = nullptr;
    check_cuda(cudaMallocAsync(reinterpret_cast<void**>(&dev), fb + 4096 + scratch224, s), "crop scratch");
    check_cuda(cudaMemcpyAsync(dev, faces.data(), faces.size() * sizeof(CropFace), cudaMemcpyHostToDevice, s), "crop faces upload");
    check_cuda(cudaMemcpyAsync(dev + fb, tables, tb, cudaMemcpyHostToDevice, s), "crop tables upload");
    uint8_t* c224 = crops224 ? crops224 : reinterpret_cast<uint8_t*>(dev + fb + 4096);
    const char* err = crop_faces_launch(frames, H, W, reinterpret_cast<const CropFace*>(dev),
                                        n, reinterpret_cast<const short*>(dev + fb), c224, crops256, s);
    cudaFreeAsync(dev, s);
    check_msg(err, "crop faces");
  });
}

int ghost_face_masks(const float* landmarks, const int* params, int n, void* masks_out, int mask_f64, void* stream) {
  return guarded([&] {
    require(landmarks && params && masks_out && n > 0, "bad argument");
    cudaStream_t s = S(stream);
    std::vector<MaskJob> jobs(n);
    for (int i = 0; i < n; ++i)
      check_msg(mask_prepare_job(&jobs[i], landmarks + static_cast<size_t>(i) * 212, params[3 * i], params[3 * i + 1],
                                 params[3 * i + 2]), "face mask");
    // own allocation (stream-ordered), so the call stays re-entrant next to ghost_paste_back's scratch
    MaskJob* dev = nullptr;
    check_cuda(cudaMallocAsync(reinterpret_cast<void**>(&dev), jobs.size() * sizeof(MaskJob), s), "mask jobs alloc");
    check_cuda(cudaMemcpyAsync(dev, jobs.data(), jobs.size() * sizeof(MaskJob), cudaMemcpyHostToDevice, s), "mask jobs upload");
    // pageable source: the copy has been staged by the time cudaMemcpyAsync returns, `jobs` may go out of scope
    const char* err = face_masks_launch(dev, n, masks_out, mask_f64, s);
    cudaFreeAsync(dev, s);
    check_msg(err, "face masks");
  });
}

int ghost_paste_back(const ghost_paste_desc* d, void* stream) {
  return guarded([&] {
    require(d, "null descriptor");
    require(d->frames_in && d->frames_out && d->n_frames > 0 && d->H > 0 && d->W > 0, "bad frame batch");
    require(d->n_faces >= 0, "bad face count");
    require(d->n_faces == 0 || (d->crops && d->masks && d->tfm && d->face_frame && d->face_crop && d->face_mask),
            "null face arrays");
    require(d->mode == GHOST_PASTE_EXACT || d->mode == GHOST_PASTE_FLOAT, "bad paste mode");
    cudaStream_t s = S(stream);
    std::vector<PasteFace> faces(d->n_faces > 0 ? d->n_faces : 1);
    std::vector<int> start(d->n_frames + 1, 0);
    int prev = 0;
    for (int i = 0; i < d->n_faces; ++i) {
      const int f = d->face_frame[i];
      require(f >= prev && f < d->n_frames, "face_frame must be non-decreasing and inside the frame batch");
      require(d->face_crop[i] >= 0 && d->face_crop[i] < d->n_crops, "face_crop out of range");
      require(d->face_mask[i] >= 0 && d->face_mask[i] < d->n_masks, "face_mask out of range");
      prev = f;
      start[f + 1]++;
      paste_prepare_face(&faces[i], d->tfm + 6 * i, d->H, d->W, d->face_crop[i], d->face_mask[i]);
      faces[i].frame = f;
    }
    for (int f = 0; f < d->n_frames; ++f) start[f + 1] += start[f];
    const size_t fb = faces.size() * sizeof(PasteFace), sb = start.size() * sizeof(int);
    const size_t fb_al = (fb + 255) / 256 * 256;
    const size_t sb_al = (sb + 255) / 256 * 256;
    char* scratch = static_cast<char*>(stream_scratch(fb_al + sb_al + paste_rz_bytes(d->n_faces > 0 ? d->n_crops : 0), s));
    check_cuda(cudaMemcpyAsync(scratch, faces.data(), fb, cudaMemcpyHostToDevice, s), "faces upload");
    check_cuda(cudaMemcpyAsync(scratch + fb_al, start.data(), sb, cudaMemcpyHostToDevice, s), "face index upload");
    DeviceTables& t = resize_tables();
    PasteParams p;
    memset(&p, 0, sizeof(p));
    p.crops = d->crops;
    p.crop_kind = d->crop_kind;
    p.masks = d->masks;
    p.mask_f64 = d->mask_f64;
    p.frames_in = d->frames_in;
    p.frames_out = d->frames_out;
    p.F = d->n_frames; p.H = d->H; p.W = d->W;
    p.faces = reinterpret_cast<const PasteFace*>(scratch);
    p.face_start = reinterpret_cast<const int*>(scratch + fb_al);
    p.rs_idx = t.idx; p.rs_a0 = t.a0; p.rs_a1 = t.a1;
    p.mode = d->mode;
    p.n_faces = d->n_faces;
    p.n_crops = d->n_crops;
    p.rz = reinterpret_cast<uint8_t*>(scratch + fb_al + sb_al);
    const char* err = paste_back_launch(p, faces.data(), s);
    cudaFreeAsync(scratch, s);  // stream-ordered: released behind the kernels that read it
    check_msg(err, "paste_back");
  });
}

/* Face boxes of a paste-back and box-only transfers of the frames (the compositing touches nothing else). */
int ghost_paste_boxes(const double* tfm, int n_faces, int H, int W, int* boxes) {
  return guarded([&] {
    require(n_faces >= 0 && H > 0 && W > 0, "bad arguments");
    require(n_faces == 0 || (tfm && boxes), "null arrays");
    for (int i = 0; i < n_faces; ++i) {
      PasteFace f;
      paste_prepare_face(&f, tfm + 6 * i, H, W, 0, 0);
      const bool empty = f.x1 <= f.x0 || f.y1 <= f.y0;
      boxes[4 * i + 0] = empty ? 0 : f.x0;
      boxes[4 * i + 1] = empty ? 0 : f.y0;
      boxes[4 * i + 2] = empty ? 0 : f.x1;
      boxes[4 * i + 3] = empty ? 0 : f.y1;
    }
  });
}

int ghost_copy_boxes(void* dst, const void* src, int n_frames, int H, int W, const int* face_frame, const int* boxes,
                     int n_faces, int to_device, void* stream) {
  return guarded([&] {
    require(dst && src && n_frames > 0 && H > 0 && W > 0 && n_faces >= 0, "bad arguments");
    require(n_faces == 0 || (face_frame && boxes), "null arrays");
    cudaStream_t s = S(stream);
    const size_t pitch = static_cast<size_t>(W) * 3;
    for (int i = 0; i < n_faces; ++i) {
      const int f = face_frame[i], x0 = boxes[4 * i], y0 = boxes[4 * i + 1], x1 = boxes[4 * i + 2], y1 = boxes[4 * i + 3];
      require(f >= 0 && f < n_frames, "face_frame out of range");
      if (x1 <= x0 || y1 <= y0) continue;
      require(x0 >= 0 && y0 >= 0 && x1 <= W && y1 <= H, "box outside the frame");
      const size_t off = (static_cast<size_t>(f) * H + y0) * pitch + static_cast<size_t>(x0) * 3;
      check_cuda(cudaMemcpy2DAsync(static_cast<uint8_t*>(dst) + off, pitch, static_cast<const uint8_t*>(src) + off, pitch,
                                   static_cast<size_t>(x1 - x0) * 3, y1 - y0,
                                   to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, s),
                 "box copy");
    }
  });
}

}  // extern "C"
