// Identity matching (utils/inference/video_processing.py:126-148).
#pragma once
#include <cuda_runtime.h>

namespace ghost {

// faces [n_faces][dim], targets [n_targets][dim] fp32 on the device (un-normalised ArcFace embeddings);
// best_idx / best_sim / matched: [n_targets] on the device.
const char* match_identities_launch(const float* faces, int n_faces, const float* targets, int n_targets, int dim,
                                    float th, int* best_idx, float* best_sim, int* matched, cudaStream_t s);

}  // namespace ghost
