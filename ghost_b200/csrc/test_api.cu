// Single-op entry points for the kernel-level parity tests.  They run one lowered op
// synchronously with weights supplied in the reference's own layouts.
#include <stdlib.h>
#include <string.h>

#include "../../include/ghost_b200.h"
#include "net.cuh"
#include "mask.cuh"

using namespace ghost;

namespace {
thread_local std::string g_err;
template <typename F>
int guarded_t(F&& f) {
  try {
    f();
    check_device_errors();
    return 0;
  } catch (const GhostError& e) {
    g_err = e.what();
    return e.code;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}
}  // namespace

extern "C" {

const char* ghost_test_last_error(void) { return g_err.c_str(); }

// GHOST_TEST_CONV_REPS=n: ghost_test_conv re-launches its (first) op n more times between two CUDA events; the mean
// duration of those launches is what ghost_test_last_ms returns (tools/conv_sweep.py)
static float g_last_ms = 0.f;
float ghost_test_last_ms(void) { return g_last_ms; }

// conv weights: kind 0 -> [Cout][Cin][k][k] (nn.Conv2d), kind 1 -> [Cin][Cout][4][4] (nn.ConvTranspose2d k4 s2 p1)
int ghost_test_conv(const void* x, int N, int H, int W, int Cin, const float* w_host, const float* bias_host, int Cout,
                    int kind, int k, int stride, int pad, int act, float slope, void* y, int mode, int bn_hint,
                    void* stream) {
  return guarded_t([&] {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int dt = mode == GHOST_F32 ? kF32 : (mode == GHOST_F16 ? kF16 : kBF16);
    DevPool pool;
    Act in;
    in.p = const_cast<void*>(x);
    in.N = N; in.H = H; in.W = W; in.C = Cin; in.dtype = dt;
    Epilogue e;
    e.act1 = act;
    e.slope = slope;
    std::vector<float> bias;
    if (bias_host) bias.assign(bias_host, bias_host + Cout);
    auto make_w = [&](const std::vector<float>& rows, int K) {
      PackedW pw;
      pw.Cout = Cout;
      pw.K = K;
      pw.rows_pad = (Cout + 255) / 256 * 256;
      if (dt == kF32 || Cin % 32) {
        pw.wf = pool.upload(rows);
      } else {
        std::vector<float> padded(static_cast<size_t>(pw.rows_pad) * K, 0.f);
        memcpy(padded.data(), rows.data(), rows.size() * sizeof(float));
        pw.wb = pool.upload_16(padded, dt == kF16);
        if (kind == 0 && k == 3 && Cout <= 3 && Cin % 64 == 0 && Cin / 64 <= kTapnMaxChunks)
          pw.wt = pool.upload_16(tapn_pack_rows(rows.data(), Cout, Cin), dt == kF16);
      }
      if (bias_host) {
        std::vector<float> b(pw.rows_pad, 0.f);
        memcpy(b.data(), bias.data(), sizeof(float) * Cout);
        pw.bias = pool.upload(b);
      }
      return pw;
    };
    if (kind == 0) {
      const int T = k * k;
      std::vector<float> rows(static_cast<size_t>(Cout) * T * Cin);
      for (int co = 0; co < Cout; ++co)
        for (int ci = 0; ci < Cin; ++ci)
          for (int t = 0; t < T; ++t)
            rows[(static_cast<size_t>(co) * T + t) * Cin + ci] = w_host[(static_cast<size_t>(co) * Cin + ci) * T + t];
      const PackedW pw = make_w(rows, T * Cin);
      const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
      Act out;
      out.p = y; out.N = N; out.H = Ho; out.W = Wo; out.C = Cout; out.dtype = dt;
      LaunchOp op;
      check_msg(lower_conv(&op, in, 0, Cin, Taps::square(k, pad, stride), Ho, Wo, pw, OutView::dense(out), e,
                           dt == kF32, device_err_word(), nullptr, bn_hint),
                "test conv lower");
      if (op.ws_bytes) op.set_workspace(pool.alloc(op.ws_bytes));
      check_msg(op.launch(s), "test conv launch");
      if (const char* ev = getenv("GHOST_TEST_CONV_REPS")) {
        const int reps = atoi(ev);
        if (reps > 0) {
          cudaEvent_t e0, e1;
          check_cuda(cudaEventCreate(&e0), "event");
          check_cuda(cudaEventCreate(&e1), "event");
          for (int i = 0; i < 3; ++i) check_msg(op.launch(s), "test conv launch");
          check_cuda(cudaEventRecord(e0, s), "event");
          for (int i = 0; i < reps; ++i) check_msg(op.launch(s), "test conv launch");
          check_cuda(cudaEventRecord(e1, s), "event");
          check_cuda(cudaEventSynchronize(e1), "event");
          float ms = 0.f;
          cudaEventElapsedTime(&ms, e0, e1);
          g_last_ms = ms / reps;
          cudaEventDestroy(e0);
          cudaEventDestroy(e1);
        }
      }
    } else {
      static const int kk[2][2] = {{1, 3}, {0, 2}};
      Act out;
      out.p = y; out.N = N; out.H = 2 * H; out.W = 2 * W; out.C = Cout; out.dtype = dt;
      const bool merged = dt != kF32 && Cin % 64 == 0 && Cout % 16 == 0;   // as the network plans do
      std::vector<float> all_rows;
      DeconvPhases dp;
      for (int ph = 0; ph < 4; ++ph) {
        const int r = ph >> 1, q = ph & 1;
        std::vector<float> rows(static_cast<size_t>(Cout) * 4 * Cin);
        for (int co = 0; co < Cout; ++co)
          for (int th = 0; th < 2; ++th)
            for (int tw = 0; tw < 2; ++tw)
              for (int ci = 0; ci < Cin; ++ci)
                rows[(static_cast<size_t>(co) * 4 + th * 2 + tw) * Cin + ci] =
                    w_host[((static_cast<size_t>(ci) * Cout + co) * 4 + kk[r][th]) * 4 + kk[q][tw]];
        if (merged) {
          all_rows.insert(all_rows.end(), rows.begin(), rows.end());
          dp.taps[ph] = Taps::deconv_phase(r, q);
          dp.o_off[ph] = OutView::deconv_phase(out, r, q, 0).off;
          continue;
        }
        const PackedW pw = make_w(rows, 4 * Cin);
        LaunchOp op;
        check_msg(lower_conv(&op, in, 0, Cin, Taps::deconv_phase(r, q), H, W, pw, OutView::deconv_phase(out, r, q, 0),
                             e, dt == kF32, device_err_word(), nullptr, bn_hint),
                  "test deconv lower");
        if (op.ws_bytes) op.set_workspace(pool.alloc(op.ws_bytes));
        check_msg(op.launch(s), "test deconv launch");
      }
      if (merged) {
        // the four phases as one launch: matrices (and bias) stacked along the rows
        PackedW pw;
        pw.Cout = 4 * Cout;
        pw.K = 4 * Cin;
        pw.rows_pad = (4 * Cout + 255) / 256 * 256;
        std::vector<float> padded(static_cast<size_t>(pw.rows_pad) * pw.K, 0.f);
        memcpy(padded.data(), all_rows.data(), all_rows.size() * sizeof(float));
        pw.wb = pool.upload_16(padded, dt == kF16);
        if (bias_host) {
          std::vector<float> b(pw.rows_pad, 0.f);
          for (int ph = 0; ph < 4; ++ph) memcpy(b.data() + ph * Cout, bias.data(), sizeof(float) * Cout);
          pw.bias = pool.upload(b);
        }
        LaunchOp op;
        check_msg(lower_conv(&op, in, 0, Cin, dp.taps[0], H, W, pw, OutView::deconv_phase(out, 0, 0, 0), e, false,
                             device_err_word(), nullptr, bn_hint, &dp),
                  "test deconv lower (merged phases)");
        if (op.ws_bytes) op.set_workspace(pool.alloc(op.ws_bytes));
        check_msg(op.launch(s), "test deconv launch (merged phases)");
      }
    }
    check_cuda(cudaStreamSynchronize(s), "test conv sync");
  });
}

int ghost_test_aad(const void* h, const void* zattr, const float* zid_host, int N, int H, int W, int c_x, int c_attr,
                   int c_id, const float* conv1_w, const float* conv1_b, const float* conv2_w, const float* conv2_b,
                   const float* fc1_w, const float* fc1_b, const float* fc2_w, const float* fc2_b,
                   const float* convh_w, float convh_b, void* out, int mode, void* stream) {
  return guarded_t([&] {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int dt = mode == GHOST_F32 ? kF32 : (mode == GHOST_F16 ? kF16 : kBF16);
    const bool bf = is16(dt);
    DevPool pool;
    const int HW = H * W;
    // identity branch on the host (fp32): gamma_id / beta_id = fc(z_id)
    std::vector<float> zfc(static_cast<size_t>(N) * 2 * c_x);
    for (int n = 0; n < N; ++n)
      for (int c = 0; c < c_x; ++c) {
        double g = fc1_b[c], b = fc2_b[c];
        for (int i = 0; i < c_id; ++i) {
          g += static_cast<double>(fc1_w[static_cast<size_t>(c) * c_id + i]) * zid_host[static_cast<size_t>(n) * c_id + i];
          b += static_cast<double>(fc2_w[static_cast<size_t>(c) * c_id + i]) * zid_host[static_cast<size_t>(n) * c_id + i];
        }
        zfc[static_cast<size_t>(n) * 2 * c_x + c] = static_cast<float>(g);
        zfc[static_cast<size_t>(n) * 2 * c_x + c_x + c] = static_cast<float>(b);
      }
    float* d_zfc = pool.upload(zfc);
    const int parts = in_stats_parts(N, HW, c_x);
    float* sums = pool.alloc_n<float>(static_cast<size_t>(N) * (2 * parts + 1) * c_x);
    float4* ncpar = pool.alloc_n<float4>(static_cast<size_t>(N) * c_x);
    float* whf = pool.alloc_n<float>(static_cast<size_t>(N) * c_x);
    float* bhf = pool.alloc_n<float>(N);
    float* d_wh = pool.upload(std::vector<float>(convh_w, convh_w + c_x));
    check_msg(in_stats_launch(h, dt, N, HW, c_x, sums, s), "stats");
    float* fsc = pool.alloc_n<float>(aad_fold_scratch_floats(N, c_x));
    check_msg(aad_fold_launch(sums, parts, N, HW, c_x, d_zfc, 2 * c_x, 0, c_x, d_wh, convh_b, ncpar, whf, bhf, fsc, s), "fold");
    const int cn = bf ? (c_x < 128 ? c_x : 128) : c_x;
    std::vector<float> rows(static_cast<size_t>(2) * c_x * c_attr), bias(2 * c_x);
    for (int c = 0; c < c_x; ++c) {
      const int t = c / cn, j = c % cn;
      const size_t rg = static_cast<size_t>(t) * 2 * cn + j, rb = rg + cn;
      memcpy(&rows[rg * c_attr], conv1_w + static_cast<size_t>(c) * c_attr, sizeof(float) * c_attr);
      memcpy(&rows[rb * c_attr], conv2_w + static_cast<size_t>(c) * c_attr, sizeof(float) * c_attr);
      bias[rg] = conv1_b[c];
      bias[rb] = conv2_b[c];
    }
    PackedW pw;
    pw.Cout = 2 * c_x;
    pw.K = c_attr;
    pw.rows_pad = (2 * c_x + 255) / 256 * 256;
    std::vector<float> bpad(pw.rows_pad, 0.f);
    memcpy(bpad.data(), bias.data(), sizeof(float) * 2 * c_x);
    pw.bias = pool.upload(bpad);
    Act za;
    za.p = const_cast<void*>(zattr);
    za.N = N; za.H = H; za.W = W; za.C = c_attr; za.dtype = dt;
    Act o;
    o.p = out; o.N = N; o.H = H; o.W = W; o.C = c_x; o.dtype = dt;
    if (bf) {
      std::vector<float> padded(static_cast<size_t>(pw.rows_pad) * c_attr, 0.f);
      memcpy(padded.data(), rows.data(), rows.size() * sizeof(float));
      pw.wb = pool.upload_16(padded, dt == kF16);
      AadArgs args;
      args.h = h; args.c_x = c_x; args.ncpar = ncpar; args.wh_fold = whf; args.bh_fold = bhf;
      if (c_x > 128) {
        float* mask = pool.alloc_n<float>(static_cast<size_t>(N) * HW);
        check_msg(aad_mask_launch(h, N, HW, c_x, whf, bhf, mask, s, dt), "aad mask");
        args.mask = mask;
      }
      LaunchOp op;
      check_msg(lower_conv(&op, za, 0, c_attr, Taps::point(), H, W, pw, OutView::dense(o), Epilogue(), false,
                           device_err_word(), &args),
                "test aad lower");
      check_msg(op.launch(s), "test aad launch");
    } else {
      pw.wf = pool.upload(rows);
      Act gb;
      gb.N = N; gb.H = H; gb.W = W; gb.C = 2 * c_x; gb.dtype = dt;
      gb.p = pool.alloc(gb.bytes());
      LaunchOp op;
      check_msg(lower_conv(&op, za, 0, c_attr, Taps::point(), H, W, pw, OutView::dense(gb), Epilogue(), true,
                           device_err_word(), nullptr),
                "test aad lower");
      check_msg(op.launch(s), "test aad launch");
      check_msg(aad_apply_launch(h, gb.p, out, dt, N, HW, c_x, ncpar, whf, bhf, s), "aad apply");
    }
    check_cuda(cudaStreamSynchronize(s), "test aad sync");
  });
}

int ghost_test_upsample2x(const void* x, void* y, int dtype, int N, int H, int W, int C, void* stream) {
  return guarded_t([&] {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    check_msg(upsample2x_launch(x, y, dtype == GHOST_F32 ? kF32 : (dtype == GHOST_F16 ? kF16 : kBF16), N, H, W, C, C, 0, s), "upsample2x");
    check_cuda(cudaStreamSynchronize(s), "upsample sync");
  });
}

// Fused upsample + InstanceNorm statistics (K5b): also folds the per-part sums on the host (in double, part order)
// into mean[N*C] / biased var[N*C] so that the caller can compare them with the statistics of y itself.
int ghost_test_upsample2x_stats(const void* x, void* y, int dtype, int N, int H, int W, int C, float* mean_host,
                                float* var_host, void* stream) {
  return guarded_t([&] {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    DevPool pool;
    const int HWo = 4 * H * W;
    const int P = upsample2x_stats_parts(N, H, W, C);
    const size_t nparts = static_cast<size_t>(N) * (2 * P + 1) * C;
    float* parts = pool.alloc_n<float>(nparts);
    check_msg(upsample2x_stats_launch(x, y, dtype == GHOST_F32 ? kF32 : (dtype == GHOST_F16 ? kF16 : kBF16), N, H, W, C, parts, s), "upsample2x_stats");
    check_cuda(cudaStreamSynchronize(s), "upsample sync");
    std::vector<float> hp(nparts);
    check_cuda(cudaMemcpy(hp.data(), parts, nparts * sizeof(float), cudaMemcpyDeviceToHost), "parts download");
    for (int n = 0; n < N; ++n)
      for (int c = 0; c < C; ++c) {
        const float* pn = hp.data() + static_cast<size_t>(n) * (2 * P + 1) * C;
        double su = 0.0, sq = 0.0;
        for (int q = 0; q < P; ++q) {
          su += pn[(2 * q) * C + c];
          sq += pn[(2 * q + 1) * C + c];
        }
        const double d = su / HWo;
        mean_host[static_cast<size_t>(n) * C + c] = static_cast<float>(pn[2 * P * C + c] + d);
        var_host[static_cast<size_t>(n) * C + c] = static_cast<float>(sq / HWo - d * d);
      }
  });
}

// Host half of the mask synthesis (no device work): the rasterised hull bitmap [224*7] and the two fixed-point
// Gaussian kernels of one face; nxny receives {nx, ny}.
int ghost_test_mask_job(const float* landmarks, int erode, int sigma_x, int sigma_y, uint32_t* bitmap, uint16_t* kx,
                        uint16_t* ky, int* nxny) {
  try {
    std::vector<MaskJob> job(1);
    const char* err = mask_prepare_job(&job[0], landmarks, erode, sigma_x, sigma_y);
    if (err) {
      g_err = err;
      return -1;
    }
    memcpy(bitmap, job[0].bitmap, sizeof(job[0].bitmap));
    memcpy(kx, job[0].kx, sizeof(job[0].kx));
    memcpy(ky, job[0].ky, sizeof(job[0].ky));
    nxny[0] = job[0].nx;
    nxny[1] = job[0].ny;
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

}  // extern "C"
