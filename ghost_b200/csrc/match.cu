// Identity matching for multi-face video (utils/inference/video_processing.py:126-148):
//   face_embeds = F.normalize(netArc(faces)); target_embeds = F.normalize(target_embeds)
//   similarity = face_embeds @ target_embeds.T; best_idxs = similarity.argmax(0)
//   matched[j] = similarity[best_idxs[j]][j] > similarity_th
// One block per target identity; fp32 like the reference (F.normalize: x / max(||x||_2, 1e-12)); argmax keeps the
// first maximum, as torch does.
#include "match.cuh"

namespace ghost {

namespace {

inline const char* last_err() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

__device__ __forceinline__ float warp_sum(float v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(256) match_kernel(const float* __restrict__ faces, int n_faces,
                                                     const float* __restrict__ targets, int dim, float th,
                                                     int* __restrict__ best_idx, float* __restrict__ best_sim,
                                                     int* __restrict__ matched) {
  extern __shared__ float tn[];  // normalised target [dim]
  __shared__ float red[8];
  __shared__ float wbest[8];
  __shared__ int widx[8];
  const int j = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* t = targets + static_cast<long long>(j) * dim;
  float s = 0.f;
  for (int k = threadIdx.x; k < dim; k += 256) s = fmaf(t[k], t[k], s);
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  float nt = 0.f;
  for (int w = 0; w < 8; ++w) nt += red[w];
  const float inv_t = 1.0f / fmaxf(sqrtf(nt), 1e-12f);
  for (int k = threadIdx.x; k < dim; k += 256) tn[k] = t[k] * inv_t;
  __syncthreads();
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = warp; i < n_faces; i += 8) {
    const float* f = faces + static_cast<long long>(i) * dim;
    float nf = 0.f;
    for (int k = lane; k < dim; k += 32) nf = fmaf(f[k], f[k], nf);
    nf = warp_sum(nf);
    const float inv_f = 1.0f / fmaxf(sqrtf(nf), 1e-12f);
    float d = 0.f;
    for (int k = lane; k < dim; k += 32) d = fmaf(f[k] * inv_f, tn[k], d);
    d = warp_sum(d);
    if (d > best) {  // faces visited in increasing order per warp: the first maximum stays
      best = d;
      bi = i;
    }
  }
  if (lane == 0) {
    wbest[warp] = best;
    widx[warp] = bi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float b = -INFINITY;
    int idx = 0x7fffffff;
    for (int w = 0; w < 8; ++w)
      if (wbest[w] > b || (wbest[w] == b && widx[w] < idx)) {
        b = wbest[w];
        idx = widx[w];
      }
    best_idx[j] = n_faces > 0 ? idx : -1;
    best_sim[j] = b;
    matched[j] = (n_faces > 0 && b > th) ? 1 : 0;
  }
}

}  // namespace

const char* match_identities_launch(const float* faces, int n_faces, const float* targets, int n_targets, int dim,
                                    float th, int* best_idx, float* best_sim, int* matched, cudaStream_t s) {
  if (n_targets <= 0) return nullptr;
  if (dim <= 0 || dim > 8192) return "match_identities: embedding dimension must be in [1, 8192]";
  match_kernel<<<n_targets, 256, dim * sizeof(float), s>>>(faces, n_faces, targets, dim, th, best_idx, best_sim, matched);
  return last_err();
}

}  // namespace ghost
