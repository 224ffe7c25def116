// Library-level entry points: version / status strings / device check and the
// host-buffer form of compute_heatmaps_from_vectors.
#include "ca_common.cuh"

#include <vector>

namespace ca { int g_last_cuda_error = 0; }

extern "C" const char* ca_version(void) { return "conceptattention_b200 0.1.0 (sm_100a)"; }

extern "C" const char* ca_status_string(ca_status s) {
  switch (s) {
    case CA_OK: return "ok";
    case CA_ERR_INVALID_ARG: return "invalid argument";
    case CA_ERR_UNSUPPORTED: return "unsupported shape";
    case CA_ERR_ALIGNMENT: return "pointer or stride not 16-byte aligned";
    case CA_ERR_ARCH: return "device is not sm_100 (B200)";
    case CA_ERR_CUDA: return "CUDA error (see ca_last_cuda_error)";
    case CA_ERR_SCRATCH_LIMIT: return "scratch need exceeds the static no-alloc array (CA_K2_NOALLOC_MAX_BYTES)";
    default: return "unknown status";
  }
}

extern "C" int ca_last_cuda_error(void) { return ca::g_last_cuda_error; }

extern "C" ca_status ca_device_check(int device) {
  cudaDeviceProp prop;
  CA_CUDA_OK(cudaGetDeviceProperties(&prop, device));
  return prop.major == 10 ? CA_OK : CA_ERR_ARCH;
}

// concept_attention_pipeline.py:29-91 with host buffers on both sides.
extern "C" ca_status ca_heatmaps_from_vectors_host(const void* image_host, const void* concept_host,
                                                   int Nt, int L, int B, int H, int P, int C, int d,
                                                   int per_head, const int* timesteps,
                                                   int n_timesteps, const int* layers, int n_layers,
                                                   int norm, int round_bf16, float* heat_host) {
  if (!image_host || !concept_host || !timesteps || !layers || !heat_host) return CA_ERR_INVALID_ARG;
  if (Nt <= 0 || L <= 0 || B <= 0 || H <= 0 || P <= 0 || C <= 0 || d <= 0 || n_timesteps <= 0 ||
      n_layers <= 0)
    return CA_ERR_INVALID_ARG;
  for (int i = 0; i < n_timesteps; ++i) if (timesteps[i] < 0 || timesteps[i] >= Nt) return CA_ERR_INVALID_ARG;
  for (int i = 0; i < n_layers; ++i) if (layers[i] < 0 || layers[i] >= L) return CA_ERR_INVALID_ARG;
  const int64_t K = (int64_t)H * d;
  const size_t img_bytes = (size_t)B * P * K * 2, con_bytes = (size_t)B * C * K * 2;
  const size_t heat_bytes = (size_t)B * C * P * sizeof(float);
  // two streams: `sc` copies pair i+1 into one buffer while `st` runs the kernel of pair i on the
  // other; copied[s] / consumed[s] order the two users of buffer s
  cudaStream_t st = nullptr, sc = nullptr;
  cudaEvent_t copied[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr};
  void *dimg[2] = {nullptr, nullptr}, *dcon[2] = {nullptr, nullptr};
  float* dacc = nullptr;
  ca_status rc = CA_OK;
  auto cleanup = [&]() {
    if (st) cudaStreamSynchronize(st);
    if (sc) cudaStreamSynchronize(sc);
    for (int i = 0; i < 2; ++i) {
      cudaFree(dimg[i]); cudaFree(dcon[i]);
      if (copied[i]) cudaEventDestroy(copied[i]);
      if (consumed[i]) cudaEventDestroy(consumed[i]);
    }
    cudaFree(dacc);
    if (st) cudaStreamDestroy(st);
    if (sc) cudaStreamDestroy(sc);
  };
#define CA_TRY(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { cleanup(); return ca::cuda_fail(_e); } } while (0)
  CA_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  CA_TRY(cudaStreamCreateWithFlags(&sc, cudaStreamNonBlocking));
  for (int i = 0; i < 2; ++i) {
    CA_TRY(cudaMalloc(&dimg[i], img_bytes)); CA_TRY(cudaMalloc(&dcon[i], con_bytes));
    CA_TRY(cudaEventCreateWithFlags(&copied[i], cudaEventDisableTiming));
    CA_TRY(cudaEventCreateWithFlags(&consumed[i], cudaEventDisableTiming));
  }
  CA_TRY(cudaMalloc(&dacc, heat_bytes));
  CA_TRY(cudaMemsetAsync(dacc, 0, heat_bytes, st));
  int it = 0;
  for (int ti = 0; ti < n_timesteps && rc == CA_OK; ++ti) {
    for (int li = 0; li < n_layers && rc == CA_OK; ++li, ++it) {
      const int64_t tl = (int64_t)timesteps[ti] * L + layers[li];
      const int s = it & 1;
      if (it >= 2) CA_TRY(cudaStreamWaitEvent(sc, consumed[s], 0));   // the kernel of pair it-2 has read buffer s
      CA_TRY(cudaMemcpyAsync(dimg[s], (const char*)image_host + tl * img_bytes, img_bytes,
                             cudaMemcpyHostToDevice, sc));
      CA_TRY(cudaMemcpyAsync(dcon[s], (const char*)concept_host + tl * con_bytes, con_bytes,
                             cudaMemcpyHostToDevice, sc));
      CA_TRY(cudaEventRecord(copied[s], sc));
      CA_TRY(cudaStreamWaitEvent(st, copied[s], 0));
      if (per_head)      // [B,H,P,d] / [B,H,C,d]
        rc = ca_heatmap_accumulate(dimg[s], (int64_t)H * P * d, (int64_t)P * d, d, dcon[s],
                                   (int64_t)H * C * d, (int64_t)C * d, d, H, d, B, P, C, norm,
                                   round_bf16, dacc, st);
      else               // [B,P,K] / [B,C,K]
        rc = ca_heatmap_accumulate(dimg[s], (int64_t)P * K, 0, K, dcon[s], (int64_t)C * K, 0, K, 1,
                                   (int)K, B, P, C, norm, round_bf16, dacc, st);
      if (rc == CA_OK) CA_TRY(cudaEventRecord(consumed[s], st));
    }
  }
  if (rc == CA_OK)
    rc = ca_finalize_heatmaps(dacc, dacc, (int64_t)B * C * P, 1.f / (float)(n_timesteps * n_layers),
                              round_bf16, st);
  if (rc == CA_OK) {
    CA_TRY(cudaMemcpyAsync(heat_host, dacc, heat_bytes, cudaMemcpyDeviceToHost, st));
    CA_TRY(cudaStreamSynchronize(st));
  }
#undef CA_TRY
  cleanup();
  return rc;
}
