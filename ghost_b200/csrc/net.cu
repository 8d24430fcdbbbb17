#include "net.cuh"

#include <math.h>
#include <string.h>

namespace ghost {

void fail(int code, const std::string& msg) { throw GhostError(code, msg); }
void check_cuda(cudaError_t e, const char* what) {
  if (e != cudaSuccess) {
    cudaGetLastError();  // clear, so the next launch check does not report this error again
    fail(-3, std::string(what) + ": " + cudaGetErrorString(e));
  }
}
void check_msg(const char* err, const std::string& where) {
  if (err) fail(-3, where + ": " + err);
}

// ------------------------------------------------------------------ DevPool
DevPool::~DevPool() {
  for (void* p : ptrs_) cudaFree(p);
}
void* DevPool::alloc(size_t bytes) {
  if (bytes == 0) bytes = 16;
  bytes = (bytes + 255) / 256 * 256;
  void* p = nullptr;
  check_cuda(cudaMalloc(&p, bytes), "cudaMalloc");
  check_cuda(cudaMemset(p, 0, bytes), "cudaMemset");
  ptrs_.push_back(p);
  total_ += bytes;
  return p;
}
float* DevPool::upload(const std::vector<float>& v) {
  float* d = alloc_n<float>(v.size());
  check_cuda(cudaMemcpy(d, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice), "upload");
  return d;
}
static inline uint16_t f2bf(float f) {  // round-to-nearest-even, NaN preserved
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return static_cast<uint16_t>(u >> 16);
}
// fp32 -> IEEE fp16 bits, round-to-nearest-even, overflow to infinity, subnormals kept (what __float2half_rn does)
static inline uint16_t f2h(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  const uint32_t sign = (u >> 16) & 0x8000u;
  const uint32_t a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return static_cast<uint16_t>(sign | 0x7e00u);          // NaN
  if (a >= 0x477ff000u) return static_cast<uint16_t>(sign | 0x7c00u);         // rounds to >= 65520: infinity
  if (a < 0x33000001u) return static_cast<uint16_t>(sign);                    // <= 2^-25: rounds to zero
  if (a < 0x38800000u) {                                                      // subnormal half
    const int e = static_cast<int>(a >> 23);                                  // biased exponent, 102..112
    const uint32_t m = (a & 0x7fffffu) | 0x800000u;                           // 24-bit significand
    const int sh = 126 - e;                                                   // 14..24: value = m * 2^-sh half ulps
    uint32_t r = m >> sh;
    const uint32_t rem = m & ((1u << sh) - 1u), half = 1u << (sh - 1);
    if (rem > half || (rem == half && (r & 1u))) ++r;
    return static_cast<uint16_t>(sign | r);
  }
  uint32_t r = a - 0x38000000u;                                               // rebias 127 -> 15
  r += 0xfffu + ((r >> 13) & 1u);
  return static_cast<uint16_t>(sign | (r >> 13));
}
__nv_bfloat16* DevPool::upload_16(const std::vector<float>& v, bool f16) {
  std::vector<uint16_t> h(v.size());
  if (f16)
    for (size_t i = 0; i < v.size(); ++i) h[i] = f2h(v[i]);
  else
    for (size_t i = 0; i < v.size(); ++i) h[i] = f2bf(v[i]);
  __nv_bfloat16* d = alloc_n<__nv_bfloat16>(v.size());
  check_cuda(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice), "upload_16");
  return d;
}

ExecPlan::~ExecPlan() {
  if (graph) cudaGraphExecDestroy(graph);
  if (graph_enc) cudaGraphExecDestroy(graph_enc);
  if (graph_gen) cudaGraphExecDestroy(graph_gen);
}

// ------------------------------------------------------------------ Net basics
Net::Net(Kind k, std::string bb, int nb, int cid, int mode)
    : kind(k), backbone(std::move(bb)), num_blocks(nb), c_id(cid), act_dtype(mode == 0 ? kF32 : (mode == 2 ? kF16 : kBF16)) {}
Net::~Net() {}

void Net::set_weight(const std::string& key, const float* data, const int64_t* shape, int ndim) {
  HostTensor t;
  t.shape.assign(shape, shape + ndim);
  const int64_t n = t.numel();
  t.data.assign(data, data + n);
  w_[key] = std::move(t);
  finalized = false;
}

const HostTensor& Net::W(const std::string& key) const {
  auto it = w_.find(key);
  if (it == w_.end()) fail(-2, "missing state_dict entry: " + key);
  return it->second;
}

PackedW Net::pack(const std::vector<float>& rows, int Cout, int K, const std::vector<float>* bias, bool need_f32,
                  bool need_bf16) {
  PackedW p;
  p.Cout = Cout;
  p.K = K;
  p.rows_pad = (Cout + 255) / 256 * 256;
  if (need_f32) p.wf = wpool_.upload(rows);
  if (need_bf16) {
    std::vector<float> padded(static_cast<size_t>(p.rows_pad) * K, 0.f);
    memcpy(padded.data(), rows.data(), rows.size() * sizeof(float));
    p.wb = wpool_.upload_16(padded, act_dtype == kF16);
    // a 3x3 conv onto <= 3 channels (rows [co][tap][ci]): also keep the tapn.cuh layout; lower_conv picks the kernel
    if (Cout <= 3 && K % (9 * 64) == 0 && K / (9 * 64) <= kTapnMaxChunks)
      p.wt = wpool_.upload_16(tapn_pack_rows(rows.data(), Cout, K / 9), act_dtype == kF16);
  }
  if (bias) {
    std::vector<float> b(p.rows_pad, 0.f);
    memcpy(b.data(), bias->data(), bias->size() * sizeof(float));
    p.bias = wpool_.upload(b);
  }
  return p;
}

void Net::finalize() {
  if (finalized) return;
  plans_.clear();
  packed_.clear();
  vecs_.clear();
  scalars_.clear();
  aad_.clear();
  if (kind == kAEI)
    finalize_aei();
  else if (kind == kSR)
    finalize_sr();
  else
    finalize_arcface();
  check_cuda(cudaDeviceSynchronize(), "finalize");
  finalized = true;
}

ExecPlan& Net::plan(int B) {
  if (!finalized) fail(-1, "network used before ghost_net_finalize");
  if (B <= 0) fail(-1, "batch size must be positive");
  auto it = plans_.find(B);
  if (it != plans_.end()) return *it->second;
  std::unique_ptr<ExecPlan> p(new ExecPlan());
  p->B = B;
  p->err_word = device_err_word();
  if (kind == kAEI)
    build_aei(*p);
  else if (kind == kSR)
    build_sr(*p);
  else
    build_arcface(*p);
  for (auto& op : p->ops) p->flops += op.flops;
  ExecPlan& ref = *p;
  plans_[B] = std::move(p);
  return ref;
}

// Captures ops [first, last) of a plan into one CUDA graph.
void Net::capture_range(ExecPlan& p, cudaStream_t s, size_t first, size_t last, cudaGraphExec_t* out) {
  // warm every kernel once (module load, attribute setup) before capturing; the warm-up pass must not count
  // as a forward for the spectral-norm power iteration, so u is put back afterwards
  std::vector<std::vector<float>> u_saved;
  for (auto& kv : sr_sn_) {
    u_saved.emplace_back(kv.second.R);
    check_cuda(cudaMemcpyAsync(u_saved.back().data(), kv.second.u, kv.second.R * sizeof(float),
                               cudaMemcpyDeviceToHost, s), "save u");
  }
  check_cuda(cudaStreamSynchronize(s), "save u sync");
  for (size_t i = first; i < last; ++i) check_msg(p.ops[i].run(s), p.ops[i].name);
  {
    size_t i = 0;
    for (auto& kv : sr_sn_)
      check_cuda(cudaMemcpyAsync(kv.second.u, u_saved[i++].data(), kv.second.R * sizeof(float),
                                 cudaMemcpyHostToDevice, s), "restore u");
  }
  check_cuda(cudaStreamSynchronize(s), "graph warm-up");
  cudaStream_t cs;
  check_cuda(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking), "capture stream");
  cudaGraph_t g = nullptr;
  check_cuda(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal), "begin capture");
  const char* err = nullptr;
  std::string where;
  for (size_t i = first; i < last; ++i) {
    err = p.ops[i].run(cs);
    if (err) {
      where = p.ops[i].name;
      break;
    }
  }
  cudaError_t ce = cudaStreamEndCapture(cs, &g);
  if (err) {
    if (g) cudaGraphDestroy(g);
    cudaStreamDestroy(cs);
    check_msg(err, where);
  }
  check_cuda(ce, "end capture");
  check_cuda(cudaGraphInstantiate(out, g, 0), "graph instantiate");
  cudaGraphDestroy(g);
  cudaStreamDestroy(cs);
}

void Net::run_plan(ExecPlan& p, cudaStream_t s, size_t first, size_t last) {
  if (last > p.ops.size()) last = p.ops.size();
  if (use_graph) {
    // three replayable ranges: the whole forward, and - for AEI_Net - the attribute encoder and the generator on
    // their own (the encoder does not depend on z_id, so a caller may run it beside the ArcFace network)
    cudaGraphExec_t* slot = nullptr;
    bool* tried = nullptr;
    if (first == 0 && last == p.ops.size()) {
      slot = &p.graph;
      tried = &p.graph_tried;
    } else if (encoder_ops > 0 && first == 0 && last == encoder_ops) {
      slot = &p.graph_enc;
      tried = &p.enc_tried;
    } else if (encoder_ops > 0 && first == encoder_ops && last == p.ops.size()) {
      slot = &p.graph_gen;
      tried = &p.gen_tried;
    }
    if (slot) {
      if (!*slot && !*tried) {
        *tried = true;
        capture_range(p, s, first, last, slot);
      }
      if (*slot) {
        check_cuda(cudaGraphLaunch(*slot, s), "graph launch");
        return;
      }
    }
  }
  for (size_t i = first; i < last; ++i) check_msg(p.ops[i].run(s), p.ops[i].name);
}

int Net::read_sn_buffer(const std::string& key, float* dst, int cap, cudaStream_t s) {
  const size_t dot = key.rfind('.');
  if (dot == std::string::npos) fail(-1, "bad spectral-norm buffer name: " + key);
  auto it = sr_sn_.find(key.substr(0, dot));
  if (it == sr_sn_.end()) fail(-1, "no spectral-norm conv named " + key.substr(0, dot));
  const std::string leaf = key.substr(dot + 1);
  const float* src = leaf == "weight_u" ? it->second.u : leaf == "weight_v" ? it->second.v : nullptr;
  if (!src) fail(-1, "spectral-norm buffer must be weight_u or weight_v");
  const int n = leaf == "weight_u" ? it->second.R : it->second.K;
  if (dst) {
    if (cap < n) fail(-1, "destination too small for " + key);
    check_cuda(cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToHost, s), "read_sn_buffer");
    check_cuda(cudaStreamSynchronize(s), "read_sn_buffer sync");
  }
  return n;
}

void Net::profile_plan(ExecPlan& p, cudaStream_t s, int reps, std::vector<float>* ms) {
  const size_t n = p.ops.size();
  ms->assign(n, 0.f);
  std::vector<cudaEvent_t> ev(n + 1);
  for (auto& e : ev) check_cuda(cudaEventCreate(&e), "event create");
  for (int r = 0; r < reps + 1; ++r) {  // first round is a warm-up
    check_cuda(cudaEventRecord(ev[0], s), "event record");
    for (size_t i = 0; i < n; ++i) {
      check_msg(p.ops[i].run(s), p.ops[i].name);
      check_cuda(cudaEventRecord(ev[i + 1], s), "event record");
    }
    check_cuda(cudaStreamSynchronize(s), "profile sync");
    if (r == 0) continue;
    for (size_t i = 0; i < n; ++i) {
      float t = 0.f;
      check_cuda(cudaEventElapsedTime(&t, ev[i], ev[i + 1]), "event elapsed");
      (*ms)[i] += t / reps;
    }
  }
  for (auto& e : ev) cudaEventDestroy(e);
}

// One pinned, device-mapped word per process: a kernel whose pipeline wait times out sets it, the
// host reads it (without synchronising) at the entry of every API call, so a device-side failure
// is reported by the next call at the latest and by ghost_check_errors() after a synchronize.
static int* g_err_host = nullptr;
int* device_err_word() {
  if (!g_err_host) {
    check_cuda(cudaHostAlloc(reinterpret_cast<void**>(&g_err_host), sizeof(int), cudaHostAllocMapped), "err word");
    *g_err_host = 0;
  }
  int* d = nullptr;
  check_cuda(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d), g_err_host, 0), "err word map");
  return d;
}
void check_device_errors() {
  if (g_err_host && *reinterpret_cast<volatile int*>(g_err_host)) {
    *g_err_host = 0;
    fail(-4, "device-side pipeline timeout in the tcgen05 conv kernel");
  }
}

}  // namespace ghost
