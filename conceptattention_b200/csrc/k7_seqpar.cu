// K7: layout kernels of the sequence-parallel exchange.
//
// Sequence-parallel mode keeps every token-wise op (LN/modulate, the linears with FULL weights,
// gated residuals) on this rank's 1/W of the image and text rows, and switches to a head split
// only inside the attention: an all-to-all turns [all heads, local rows] into
// [local heads, all rows] for K2 and a second one turns K2's output back.  NCCL moves contiguous
// per-destination chunks; these kernels are the two permutations either side of K2 and the one after
// the return exchange.  B = 1.  Rows are 256-byte head vectors (or Hl of them): every access is a
// 16-byte vector, fully coalesced; pure HBM copies (2 x bytes moved).
//
//   local row order of a rank  : [img rows (Pl) | txt rows (Tl) | concept rows (C, replicated)]
//   K2 row order               : [img (W*Pl)    | txt (W*Tl)    | concept (C)]
#include "ca_common.cuh"

namespace ca {

// recv [W][3][Hl][Nl][128] -> q,k,v [Hl][N][128]; concept rows are taken from this rank's own chunk
__global__ void __launch_bounds__(256)
seq_unpack_qkv_kernel(const uint4* __restrict__ recv, int W, int self, int Hl, int Pl, int Tl, int C,
                      uint4* __restrict__ q, uint4* __restrict__ k, uint4* __restrict__ v) {
  const int Nl = Pl + Tl + C;
  const int64_t N = (int64_t)W * (Pl + Tl) + C;
  const int64_t total = (int64_t)W * 3 * Hl * Nl * 16;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c16 = (int)(i & 15);
  int64_t r = i >> 4;
  const int row = (int)(r % Nl); r /= Nl;
  const int hl = (int)(r % Hl); r /= Hl;
  const int t = (int)(r % 3);
  const int w = (int)(r / 3);
  int64_t dst_row;
  if (row < Pl) dst_row = (int64_t)w * Pl + row;
  else if (row < Pl + Tl) dst_row = (int64_t)W * Pl + (int64_t)w * Tl + (row - Pl);
  else if (w == self) dst_row = (int64_t)W * (Pl + Tl) + (row - Pl - Tl);
  else return;
  uint4* dst = t == 0 ? q : (t == 1 ? k : v);
  dst[((int64_t)hl * N + dst_row) * 16 + c16] = __ldg(recv + i);
}

// K2 outputs out_img [W*Pl][Hl*128], out_txtc [W*Tl + C][Hl*128] -> send [W][Nl][Hl*128]
// (every destination also gets the C concept rows of this rank's heads)
__global__ void __launch_bounds__(256)
seq_pack_out_kernel(const uint4* __restrict__ out_img, const uint4* __restrict__ out_txtc, int W,
                    int Hl, int Pl, int Tl, int C, uint4* __restrict__ send) {
  const int Nl = Pl + Tl + C, rv = Hl * 16;                 // 16-byte vectors per row
  const int64_t total = (int64_t)W * Nl * rv;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)(i % rv);
  int64_t r = i / rv;
  const int row = (int)(r % Nl);
  const int w = (int)(r / Nl);
  const uint4* src;
  if (row < Pl) src = out_img + ((int64_t)w * Pl + row) * rv;
  else if (row < Pl + Tl) src = out_txtc + ((int64_t)w * Tl + (row - Pl)) * rv;
  else src = out_txtc + ((int64_t)W * Tl + (row - Pl - Tl)) * rv;
  send[i] = __ldg(src + c);
}

// recv [W][Nl][Hl*128] (chunk w = heads of rank w) -> attn_img [Pl][W*Hl*128], attn_tc [Tl+C][W*Hl*128]
__global__ void __launch_bounds__(256)
seq_unpack_out_kernel(const uint4* __restrict__ recv, int W, int Hl, int Pl, int Tl, int C,
                      uint4* __restrict__ attn_img, uint4* __restrict__ attn_tc) {
  const int Nl = Pl + Tl + C, rv = Hl * 16;
  const int64_t total = (int64_t)W * Nl * rv;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)(i % rv);
  int64_t r = i / rv;
  const int row = (int)(r % Nl);
  const int w = (int)(r / Nl);
  uint4* dst = row < Pl ? attn_img + (int64_t)row * W * rv : attn_tc + (int64_t)(row - Pl) * W * rv;
  dst[(int64_t)w * rv + c] = __ldg(recv + i);
}

// ---- fused exchange: pull straight out of the peers' buffers over NVLink ---------------------------
// peers[w] = base of rank w's symmetric allocation as mapped into THIS process (P2P / fabric
// handles, exchanged once by the host side); every rank uses the same byte offsets inside it.
// Same permutation as seq_unpack_qkv with recv[w] := (rank w's send buffer)[self]: the all-to-all
// and its staging buffer disappear, the loads below ARE the exchange.
__global__ void __launch_bounds__(256)
seq_gather_qkv_kernel(const uint8_t* const* __restrict__ peers, int64_t send_off, int W, int self, int Hl,
                      int Pl, int Tl, int C, uint4* __restrict__ q, uint4* __restrict__ k,
                      uint4* __restrict__ v) {
  const int Nl = Pl + Tl + C;
  const int64_t N = (int64_t)W * (Pl + Tl) + C;
  const int64_t chunk = (int64_t)3 * Hl * Nl * 16;          // 16-byte vectors per destination chunk
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= chunk * W) return;
  const int w = (int)(i / chunk);
  const int64_t j = i - (int64_t)w * chunk;                 // offset inside the chunk
  const int c16 = (int)(j & 15);
  int64_t r = j >> 4;
  const int row = (int)(r % Nl); r /= Nl;
  const int hl = (int)(r % Hl);
  const int t = (int)(r / Hl);
  int64_t dst_row;
  if (row < Pl) dst_row = (int64_t)w * Pl + row;
  else if (row < Pl + Tl) dst_row = (int64_t)W * Pl + (int64_t)w * Tl + (row - Pl);
  else if (w == self) dst_row = (int64_t)W * (Pl + Tl) + (row - Pl - Tl);
  else return;
  const uint4* src = reinterpret_cast<const uint4*>(peers[w] + send_off) + (int64_t)self * chunk + j;
  uint4* dst = t == 0 ? q : (t == 1 ? k : v);
  dst[((int64_t)hl * N + dst_row) * 16 + c16] = *src;
}

// attn_img [Pl][W*Hl*128], attn_tc [Tl+C][W*Hl*128] <- rows of THIS rank out of every peer's K2
// outputs out_img [W*Pl][Hl*128] (at img_off) and out_txtc [W*Tl+C][Hl*128] (at tc_off): what
// seq_pack_out + all-to-all + seq_unpack_out did in three passes.
__global__ void __launch_bounds__(256)
seq_gather_out_kernel(const uint8_t* const* __restrict__ peers, int64_t img_off, int64_t tc_off, int W,
                      int self, int Hl, int Pl, int Tl, int C, uint4* __restrict__ attn_img,
                      uint4* __restrict__ attn_tc) {
  const int Nl = Pl + Tl + C, rv = Hl * 16;
  const int64_t total = (int64_t)W * Nl * rv;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)(i % rv);
  int64_t r = i / rv;
  const int row = (int)(r % Nl);
  const int w = (int)(r / Nl);
  const uint4* src;
  if (row < Pl) src = reinterpret_cast<const uint4*>(peers[w] + img_off) + ((int64_t)self * Pl + row) * rv;
  else if (row < Pl + Tl)
    src = reinterpret_cast<const uint4*>(peers[w] + tc_off) + ((int64_t)self * Tl + (row - Pl)) * rv;
  else src = reinterpret_cast<const uint4*>(peers[w] + tc_off) + ((int64_t)W * Tl + (row - Pl - Tl)) * rv;
  uint4* dst = row < Pl ? attn_img + (int64_t)row * W * rv : attn_tc + (int64_t)(row - Pl) * W * rv;
  dst[(int64_t)w * rv + c] = src[c];
}

static bool seq_args_ok(int W, int Hl, int Pl, int Tl, int C) {
  return W > 0 && Hl > 0 && Pl >= 0 && Tl >= 0 && C >= 0;
}

}  // namespace ca

extern "C" ca_status ca_seq_unpack_qkv(const void* recv, int W, int self, int Hl, int Pl, int Tl, int C,
                                       void* q, void* k, void* v, void* stream) {
  using namespace ca;
  if (!seq_args_ok(W, Hl, Pl, Tl, C) || self < 0 || self >= W || !recv || !q || !k || !v)
    return CA_ERR_INVALID_ARG;
  const void* ptrs[] = {recv, q, k, v};
  for (const void* p : ptrs) if (!aligned16(p)) return CA_ERR_ALIGNMENT;
  const int64_t total = (int64_t)W * 3 * Hl * (Pl + Tl + C) * 16;
  if (total == 0) return CA_OK;
  const int64_t blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
  seq_unpack_qkv_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)recv, W, self, Hl, Pl, Tl, C, (uint4*)q, (uint4*)k, (uint4*)v);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_seq_pack_out(const void* out_img, const void* out_txtc, int W, int Hl, int Pl,
                                     int Tl, int C, void* send, void* stream) {
  using namespace ca;
  if (!seq_args_ok(W, Hl, Pl, Tl, C) || !send || (Pl > 0 && !out_img) || (Tl + C > 0 && !out_txtc))
    return CA_ERR_INVALID_ARG;
  const void* ptrs[] = {out_img, out_txtc, send};
  for (const void* p : ptrs) if (p && !aligned16(p)) return CA_ERR_ALIGNMENT;
  const int64_t total = (int64_t)W * (Pl + Tl + C) * Hl * 16;
  if (total == 0) return CA_OK;
  const int64_t blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
  seq_pack_out_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)out_img, (const uint4*)out_txtc, W, Hl, Pl, Tl, C, (uint4*)send);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_seq_unpack_out(const void* recv, int W, int Hl, int Pl, int Tl, int C,
                                       void* attn_img, void* attn_tc, void* stream) {
  using namespace ca;
  if (!seq_args_ok(W, Hl, Pl, Tl, C) || !recv || (Pl > 0 && !attn_img) || (Tl + C > 0 && !attn_tc))
    return CA_ERR_INVALID_ARG;
  const void* ptrs[] = {recv, attn_img, attn_tc};
  for (const void* p : ptrs) if (p && !aligned16(p)) return CA_ERR_ALIGNMENT;
  const int64_t total = (int64_t)W * (Pl + Tl + C) * Hl * 16;
  if (total == 0) return CA_OK;
  const int64_t blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
  seq_unpack_out_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)recv, W, Hl, Pl, Tl, C, (uint4*)attn_img, (uint4*)attn_tc);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_seq_gather_qkv(const void* const* peers_dev, int64_t send_off, int W, int self,
                                       int Hl, int Pl, int Tl, int C, void* q, void* k, void* v,
                                       void* stream) {
  using namespace ca;
  if (!seq_args_ok(W, Hl, Pl, Tl, C) || self < 0 || self >= W || !peers_dev || !q || !k || !v || send_off < 0)
    return CA_ERR_INVALID_ARG;
  const void* ptrs[] = {q, k, v};
  for (const void* p : ptrs) if (!aligned16(p)) return CA_ERR_ALIGNMENT;
  if (send_off % 16) return CA_ERR_ALIGNMENT;
  const int64_t total = (int64_t)W * 3 * Hl * (Pl + Tl + C) * 16;
  if (total == 0) return CA_OK;
  const int64_t blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
  seq_gather_qkv_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const uint8_t* const*)peers_dev, send_off, W, self, Hl, Pl, Tl, C, (uint4*)q, (uint4*)k, (uint4*)v);
  CA_LAUNCH_CHECK();
  return CA_OK;
}

extern "C" ca_status ca_seq_gather_out(const void* const* peers_dev, int64_t img_off, int64_t tc_off, int W,
                                       int self, int Hl, int Pl, int Tl, int C, void* attn_img,
                                       void* attn_tc, void* stream) {
  using namespace ca;
  if (!seq_args_ok(W, Hl, Pl, Tl, C) || self < 0 || self >= W || !peers_dev || img_off < 0 || tc_off < 0 ||
      (Pl > 0 && !attn_img) || (Tl + C > 0 && !attn_tc))
    return CA_ERR_INVALID_ARG;
  const void* ptrs[] = {attn_img, attn_tc};
  for (const void* p : ptrs) if (p && !aligned16(p)) return CA_ERR_ALIGNMENT;
  if (img_off % 16 || tc_off % 16) return CA_ERR_ALIGNMENT;
  const int64_t total = (int64_t)W * (Pl + Tl + C) * Hl * 16;
  if (total == 0) return CA_OK;
  const int64_t blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffLL) return CA_ERR_UNSUPPORTED;
  seq_gather_out_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const uint8_t* const*)peers_dev, img_off, tc_off, W, self, Hl, Pl, Tl, C, (uint4*)attn_img,
      (uint4*)attn_tc);
  CA_LAUNCH_CHECK();
  return CA_OK;
}
