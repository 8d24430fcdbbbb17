// Host-side model plans: weight store, packing (BN fold, tap-major repack),
// activation buffers, the op list of one forward and its CUDA graph.
#pragma once
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "conv.cuh"
#include "paste.cuh"

namespace ghost {

struct GhostError : std::runtime_error {
  int code;
  GhostError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
[[noreturn]] void fail(int code, const std::string& msg);
void check_cuda(cudaError_t e, const char* what);
void check_msg(const char* err, const std::string& where);  // nullptr = ok, else throws GHOST_ERR_CUDA

int* device_err_word();
void check_device_errors();

struct HostTensor {
  std::vector<float> data;
  std::vector<int64_t> shape;
  int64_t numel() const {
    int64_t n = 1;
    for (auto s : shape) n *= s;
    return n;
  }
};

// Owns device allocations.
class DevPool {
 public:
  ~DevPool();
  void* alloc(size_t bytes);
  template <typename T>
  T* alloc_n(size_t n) {
    return reinterpret_cast<T*>(alloc(n * sizeof(T)));
  }
  float* upload(const std::vector<float>& v);
  // 16-bit copy of v in the net's storage type: bf16, or IEEE fp16 bits behind the same pointer type
  __nv_bfloat16* upload_16(const std::vector<float>& v, bool f16);
  __nv_bfloat16* upload_bf16(const std::vector<float>& v) { return upload_16(v, false); }
  size_t total_bytes() const { return total_; }

 private:
  std::vector<void*> ptrs_;
  size_t total_ = 0;
};

// Kernel class of an op, for the per-class roofline accounting of the bench.
enum OpClass : int {
  kOpConvTC = 0,   // tcgen05 implicit-GEMM conv / deconv phase / linear (tensor bound)
  kOpAadTC = 1,    // tcgen05 gamma/beta GEMM with the fused AAD blend epilogue (HBM bound)
  kOpConvSimt = 2, // CUDA-core conv (fp32 mode, Cin = 3 layers)
  kOpStats = 3,    // instance-norm statistics (HBM bound)
  kOpUpsample = 4, // bilinear x2 (HBM bound)
  kOpMisc = 5      // folds, scatter, per-channel affine, layout conversion
};

struct Op {
  std::string name;
  std::function<const char*(cudaStream_t)> run;
  double flops = 0.0;  // dense multiply-add flops (2*MAC) of contraction ops
  double bytes = 0.0;  // algorithmic HBM bytes of bandwidth-bound ops (SURVEY 8d accounting)
  int cls = kOpMisc;
};

// A conv's weights in both packings.
struct ConvWeights {
  PackedW pw;
};

// One executable forward for a fixed batch size.
struct ExecPlan {
  int B = 0;
  DevPool pool;
  std::vector<Op> ops;
  cudaGraphExec_t graph = nullptr;
  bool graph_tried = false;
  cudaGraphExec_t graph_enc = nullptr, graph_gen = nullptr;  // AEI_Net: attribute encoder / generator on their own
  bool enc_tried = false, gen_tried = false;
  int* err_word = nullptr;
  // boundary buffers
  Act x_in;          // NHWC-3 network input
  Act zid;           // [B,1,1,c_id]
  Act y;             // NHWC-3 generator output
  Act attr[8];       // z_attr1..8 (NHWC); for unet they are the concat buffers
  Act emb_in;        // arcface: [B,112,112,3]
  float* emb_out = nullptr;  // arcface: [B,512] fp32
  void* stage = nullptr;     // staging for broadcast z_id
  double flops = 0.0;
  std::map<std::string, Act> named;  // intermediate activations by name (block-level parity tests, debugging)
  ~ExecPlan();
};

class Net {
 public:
  enum Kind { kAEI, kArcFace, kSR };
  Net(Kind kind, std::string backbone, int num_blocks, int c_id, int mode);
  ~Net();

  void set_weight(const std::string& key, const float* data, const int64_t* shape, int ndim);
  void finalize();
  ExecPlan& plan(int B);
  void run_plan(ExecPlan& p, cudaStream_t s, size_t first_op = 0, size_t last_op = static_cast<size_t>(-1));
  void capture_range(ExecPlan& p, cudaStream_t s, size_t first, size_t last, cudaGraphExec_t* out);
  // Runs every op `reps` times bracketed by CUDA events on `s`; ms[i] = mean duration of op i.
  void profile_plan(ExecPlan& p, cudaStream_t s, int reps, std::vector<float>* ms);

  Kind kind;
  std::string backbone;
  int num_blocks, c_id;
  int act_dtype;  // kF32, or the 16-bit storage type of the tensor-core path: kBF16 / kF16
  bool use_graph = false;
  bool finalized = false;
  size_t encoder_ops = 0;  // ops [0, encoder_ops) compute the attribute maps

 private:
  const HostTensor& W(const std::string& key) const;
  bool has(const std::string& key) const { return w_.count(key) != 0; }
  // packing
  PackedW pack(const std::vector<float>& w_rows, int Cout, int K, const std::vector<float>* bias, bool need_f32,
               bool need_bf16);
  void build_aei(ExecPlan& p);
  void build_arcface(ExecPlan& p);
  void finalize_aei();
  void finalize_arcface();
  void build_sr(ExecPlan& p);
  void finalize_sr();

  std::map<std::string, HostTensor> w_;
  DevPool wpool_;
  std::map<std::string, PackedW> packed_;       // by logical layer name
  std::map<std::string, float*> vecs_;          // small fp32 vectors (conv_h weights, prelu slopes, ...)
  std::map<std::string, float> scalars_;
  std::map<int, std::unique_ptr<ExecPlan>> plans_;
  int zfc_cols_ = 0;
  struct AadInfo { std::string name; int c_x, c_attr, off_gamma, off_beta; };
  std::map<std::string, AadInfo> aad_;
  // super-resolution generator (sr.cu): spectral-norm state per conv (device; u advances with every forward, like
  // the reference's weight_u buffer in train() mode), channel slice of each SPADE's hidden activation
  struct SnState { int R = 0, K = 0; const float* W = nullptr; float *u = nullptr, *v = nullptr, *inv_sigma = nullptr, *scratch = nullptr; };
  struct SrSlice { int ri, c0, nhp; };
  std::map<std::string, SnState> sr_sn_;
  std::map<std::string, SrSlice> sr_slices_;
  int sr_actv_c_[6] = {0, 0, 0, 0, 0, 0};
  int sr_actv_need_[6] = {0, 0, 0, 0, 0, 0};  // end of the widest read window per resolution

 public:
  // copies a spectral-norm buffer ("<conv>.weight_u" / "<conv>.weight_v") to the host; returns its length
  int read_sn_buffer(const std::string& key, float* dst, int cap, cudaStream_t s);
};

}  // namespace ghost
