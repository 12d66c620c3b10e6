// Batched leaf evaluation on the GPU: battle encoding -> embedding nets -> main MLP, one fused kernel.
//
// Reference behaviour: NN::Battle::NetworkImpl::value_policy_inference (cpp/include/nn/battle/network.h:
// 102-175), sparse encoders (cpp/include/encode/battle/battle.h:26-45,74-83,155-162,181-195,208-214,
// 262-283,319-349,381-409,480-489,544-551), policy index (encode/battle/policy.h:29-58), MainNet
// (nn/battle/main-net.h:66-107), parameter file (nn/affine.h:34-69, network.h:52-70, search.cc:127-147).
//
// Kernel shape (v1, fp32 on the CUDA cores): one warp per evaluation. The 384-byte record is staged
// in shared memory; the encoder never materialises the 2 834-float dense input — lanes evaluate the
// <= 52 candidate features of a slot in parallel, a ballot compacts the non-zeros, and the first
// layer is the gather-sum over those columns the reference's sparse Affine::propagate performs
// (affine.h:86-102) with each lane owning 4 hidden units. Second layers read their (transposed)
// weights from shared memory; the 768-wide battle embedding lives only in shared memory.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "context.h"
#include "gen1_state.cuh"

using namespace oakb200;

struct NetDims {
  int act;                      // 1 relu, 2 clamp
  int phd, pod, ahd, aod;       // pokemon / active hidden and out dims
  int side, in, hd, vhd, pohd;  // side embedding, main input (2*side), main hidden, value hidden, policy hidden
};

// device parameter block: every matrix stored TRANSPOSED ([in][out]) so lanes read consecutive outputs
struct oakb200_net {
  NetDims d;
  float* blob = nullptr;  // one device allocation
  // offsets (floats) into blob
  size_t pk0T, pk0b, pk1T, pk1b, ac0T, ac0b, ac1T, ac1b, fc0T, fc0b, fc1T, fc1b, v2T, v2b, v3w, v3b, p1aT, p1ab, p1bW, p1bb, p2aT,
      p2ab, p2bW, p2bb;
  size_t floats = 0;
  // tensor-core path (value_net_tc.cuh): packed split-bf16 weight chunks, chunk program, fp32 constants
  bool tc_ok = false;
  uint8_t* tc_wpack = nullptr;
  void* tc_prog = nullptr;
  float* tc_consts = nullptr;
  int tc_nchunks = 0;
};

namespace {

struct HostAffine {
  uint32_t in = 0, out = 0;
  const float* bias = nullptr;
  const float* w = nullptr;  // [out][in]
};

const uint8_t* read_affine(const uint8_t* p, const uint8_t* end, HostAffine& a) {
  if (!p || p + 8 > end) return nullptr;
  memcpy(&a.in, p, 4);
  memcpy(&a.out, p + 4, 4);
  p += 8;
  const size_t nb = (size_t)a.out * 4, nw = (size_t)a.out * a.in * 4;
  if (a.in == 0 || a.out == 0 || a.in > 4096 || a.out > 4096 || p + nb + nw > end) return nullptr;
  a.bias = reinterpret_cast<const float*>(p);
  a.w = reinterpret_cast<const float*>(p + nb);
  return p + nb + nw;
}

size_t put_transposed(std::vector<float>& blob, const HostAffine& a) {
  const size_t off = blob.size();
  blob.resize(off + (size_t)a.in * a.out);
  for (uint32_t o = 0; o < a.out; o++)
    for (uint32_t k = 0; k < a.in; k++) {
      float v;
      memcpy(&v, reinterpret_cast<const uint8_t*>(a.w) + ((size_t)o * a.in + k) * 4, 4);
      blob[off + (size_t)k * a.out + o] = v;
    }
  while (blob.size() % 4) blob.push_back(0.0f);
  return off;
}
size_t put_raw(std::vector<float>& blob, const float* src, size_t n) {
  const size_t off = blob.size();
  blob.resize(off + n);
  memcpy(blob.data() + off, src, n * 4);
  while (blob.size() % 4) blob.push_back(0.0f);
  return off;
}

// builds the tensor-core path's packed weights / program (defined at the end of this file)
cudaError_t tc_pack(oakb200_net* n, const HostAffine* a, const NetDims& d);

}  // namespace

void oakb200_net_release(oakb200_ctx* ctx) {
  if (ctx && ctx->net) {
    if (ctx->net->blob) cudaFree(ctx->net->blob);
    if (ctx->net->tc_wpack) cudaFree(ctx->net->tc_wpack);
    if (ctx->net->tc_prog) cudaFree(ctx->net->tc_prog);
    if (ctx->net->tc_consts) cudaFree(ctx->net->tc_consts);
    delete ctx->net;
    ctx->net = nullptr;
  }
}

static int net_load_impl(oakb200_ctx* ctx, const uint8_t* bytes, size_t len, bool discrete);

extern "C" int oakb200_net_load(oakb200_ctx* ctx, const uint8_t* bytes, size_t len) { return net_load_impl(ctx, bytes, len, false); }

// "discrete" agent (reference cpp/src/search.cc:105-123,133-141): the float embedding nets feed an int8 main net. The main
// net's parameters are converted exactly as NN::Battle::Quantized::AffineTransform::try_copy_parameters does
// (quantized/affine.h:74-96: int8 = float * 64, int32 bias = float * 64 * 127, both truncated) and kept as
// integer-valued floats, so both evaluation kernels run them unchanged: every product and sum (< 2^24) is exact in fp32 /
// in the bf16 x bf16 -> fp32 tensor-core path.
extern "C" int oakb200_net_load_discrete(oakb200_ctx* ctx, const uint8_t* bytes, size_t len) {
  if (!ctx || !bytes || len < 8) return fail(ctx, OAKB200_EINVAL, "net_load_discrete: null or short parameter image");
  if (bytes[0] + 1 != 2)
    return fail(ctx, OAKB200_EINVAL, "net_load_discrete: the parsed header does not encode clamped activations");  // search.cc:137-141
  std::vector<uint8_t> q(bytes, bytes + len);
  const uint8_t *p = q.data() + 8, *end = q.data() + len;
  for (int i = 0; i < 12; i++) {
    HostAffine a;
    const uint8_t* next = read_affine(p, end, a);
    if (!next) return fail(ctx, OAKB200_EINVAL, "net_load_discrete: truncated parameter image");
    if (i >= 4) {
      uint8_t* bw = q.data() + (reinterpret_cast<const uint8_t*>(a.bias) - q.data());
      const size_t nb = a.out, nw = (size_t)a.out * a.in;
      for (size_t k = 0; k < nb + nw; k++) {
        float v;
        memcpy(&v, bw + 4 * k, 4);
        if (k >= nb && !(v < 2.0f && v > -2.0f))
          return fail(ctx, OAKB200_EINVAL, "net_load_discrete: a main-net weight is outside (-2, 2)");  // affine.h:88-95
        const float r = k < nb ? (float)static_cast<int32_t>(v * 64 * 127) : (float)static_cast<int8_t>(v * 64);
        memcpy(bw + 4 * k, &r, 4);
      }
    }
    p = next;
  }
  return net_load_impl(ctx, q.data(), len, true);
}

static int net_load_impl(oakb200_ctx* ctx, const uint8_t* bytes, size_t len, bool discrete) {
  if (!ctx || !bytes || len < 8) return fail(ctx, OAKB200_EINVAL, "net_load: null or short parameter image");
  HostAffine a[12];
  const uint8_t *p = bytes + 8, *end = bytes + len;
  for (int i = 0; i < 12; i++) p = read_affine(p, end, a[i]);
  if (p != end) return fail(ctx, OAKB200_EINVAL, "net_load: parameter image does not end at the last Affine block");
  NetDims d;
  d.act = discrete ? 3 : bytes[0] + 1;
  d.phd = a[0].out; d.pod = a[1].out; d.ahd = a[2].out; d.aod = a[3].out;
  d.side = (1 + d.aod) + 5 * (1 + d.pod);
  d.in = 2 * d.side; d.hd = a[4].out; d.vhd = a[6].out; d.pohd = a[8].out;
  const bool ok = (d.act >= 1 && d.act <= 3) && a[0].in == 198 && a[1].in == (uint32_t)d.phd && a[2].in == 427 &&
                  a[3].in == (uint32_t)d.ahd && a[4].in == (uint32_t)d.in && a[5].in == (uint32_t)d.hd && a[5].out == (uint32_t)d.hd &&
                  a[6].in == (uint32_t)d.hd && a[7].in == (uint32_t)d.vhd && a[7].out == 1 && a[8].in == (uint32_t)d.hd &&
                  a[9].in == (uint32_t)d.pohd && a[9].out == 315 && a[10].in == (uint32_t)d.hd && a[10].out == (uint32_t)d.pohd &&
                  a[11].in == (uint32_t)d.pohd && a[11].out == 315;
  if (!ok) return fail(ctx, OAKB200_EINVAL, "net_load: layer shapes are not a battle network");
  if (d.phd > 128 || d.ahd > 128 || (d.phd & 3) || (d.ahd & 3) || d.pod > 96 || d.aod > 96 || d.hd > 64 || d.vhd > 64 || d.pohd > 64)
    return fail(ctx, OAKB200_EINVAL, "net_load: dims exceed kernel limits (hidden <= 128 and a multiple of 4, embedding <= 96, main/head hidden <= 64)");
  CU(ctx, cudaSetDevice(ctx->device));
  oakb200_net_release(ctx);
  oakb200_net* n = new oakb200_net();
  n->d = d;
  std::vector<float> blob;
  n->pk0T = put_transposed(blob, a[0]);  n->pk0b = put_raw(blob, a[0].bias, a[0].out);
  n->pk1T = put_transposed(blob, a[1]);  n->pk1b = put_raw(blob, a[1].bias, a[1].out);
  n->ac0T = put_transposed(blob, a[2]);  n->ac0b = put_raw(blob, a[2].bias, a[2].out);
  n->ac1T = put_transposed(blob, a[3]);  n->ac1b = put_raw(blob, a[3].bias, a[3].out);
  n->fc0T = put_transposed(blob, a[4]);  n->fc0b = put_raw(blob, a[4].bias, a[4].out);
  n->fc1T = put_transposed(blob, a[5]);  n->fc1b = put_raw(blob, a[5].bias, a[5].out);
  n->v2T = put_transposed(blob, a[6]);   n->v2b = put_raw(blob, a[6].bias, a[6].out);
  n->v3w = put_raw(blob, a[7].w, a[7].in); n->v3b = put_raw(blob, a[7].bias, 1);
  n->p1aT = put_transposed(blob, a[8]);  n->p1ab = put_raw(blob, a[8].bias, a[8].out);
  n->p1bW = put_raw(blob, a[9].w, (size_t)315 * d.pohd);  n->p1bb = put_raw(blob, a[9].bias, 315);
  n->p2aT = put_transposed(blob, a[10]); n->p2ab = put_raw(blob, a[10].bias, a[10].out);
  n->p2bW = put_raw(blob, a[11].w, (size_t)315 * d.pohd); n->p2bb = put_raw(blob, a[11].bias, 315);
  n->floats = blob.size();
  cudaError_t e = cudaMalloc(&n->blob, blob.size() * 4);
  if (e == cudaSuccess) e = cudaMemcpy(n->blob, blob.data(), blob.size() * 4, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    if (n->blob) cudaFree(n->blob);
    delete n;
    return fail(ctx, OAKB200_ECUDA, "net_load: upload", e);
  }
  ctx->net = n;
  e = tc_pack(n, a, d);
  if (e != cudaSuccess) {
    oakb200_net_release(ctx);
    return fail(ctx, OAKB200_ECUDA, "net_load: tensor-core weight upload", e);
  }
  return OAKB200_OK;
}

// ------------------------------------------------------------------------------------------------
#include "gen1_mem_ops.cuh"

namespace {

const int EVAL_WARPS = 8;  // evaluations in flight per CTA

struct NetView {
  NetDims d;
  const float *pk0T, *pk0b, *pk1T, *pk1b, *ac0T, *ac0b, *ac1T, *ac1b, *fc0T, *fc0b, *fc1T, *fc1b, *v2T, *v2b, *v3w, *v3b, *p1aT, *p1ab,
      *p1bW, *p1bb, *p2aT, *p2ab, *p2bW, *p2bb;
};

// act: 1 relu, 2 clamp to [0, 1]; discrete mode (NetDims::act == 3) uses 3 = clamp then floor(127 x) for the embedding
// outputs (static_cast<uint8_t>(127 * f), nn/battle/cache.h:106,234) and 4 = the int8 net's ClippedReLU
// clamp(x >> 6, 0, 127) (quantized/clipped_relu.h:53-98) on exact integer-valued floats
__device__ __forceinline__ float activate(float x, int act) {
  if (act == 4) return fminf(fmaxf(floorf(x * 0.015625f), 0.0f), 127.0f);
  x = fmaxf(x, 0.0f);
  if (act >= 2) x = fminf(x, 1.0f);
  return act == 3 ? floorf(127.0f * x) : x;
}
__device__ __forceinline__ int status_index(uint32_t status, uint32_t sleeps) {  // battle.h:106-123
  if ((status & 7u) == 0) return __ffs(status) - 1 - 3;
  if (!(status & 0x80u)) return 3 + (int)sleeps;
  return 14 - (int)(status & 7u);
}

// candidate feature c (0..11) of a stored Pokemon (battle.h:208-214); off = column offset of the block
__device__ __forceinline__ bool pokemon_feature(const uint8_t* pk, uint32_t sleep, int c, int off, int& idx, float& val) {
  val = 1.0f;
  if (c < 5) {
    const uint32_t s = pk[2 * c] | (pk[2 * c + 1] << 8);
    idx = off + c;
    val = (float)s / (c == 0 ? 703.0f : 999.0f);
    return true;
  }
  if (c < 9) {
    const uint32_t id = pk[PK_MOVES + 2 * (c - 5)], pp = pk[PK_MOVES + 2 * (c - 5) + 1];
    idx = off + 5 + (int)id - 1;
    return id != MOVE_None && id != MOVE_Struggle && pp != 0;
  }
  if (c == 9) {
    const uint32_t st = pk[PK_STATUS];
    idx = off + 169 + (st ? status_index(st, sleep) : 0);
    return st != 0;
  }
  const uint32_t t1 = pk[PK_TYPES] & 15u, t2 = pk[PK_TYPES] >> 4;
  idx = off + 183 + (int)(c == 10 ? t1 : t2);
  return c == 10 || t2 != t1;
}

// candidate feature c (0..39) of the Active block (battle.h:480-489); a = 32-byte ActivePokemon
__device__ __forceinline__ bool active_feature(const uint8_t* a, uint32_t dur, int c, int& idx, float& val) {
  val = 1.0f;
  if (c < 5) {
    const uint32_t s = a[2 * c] | (a[2 * c + 1] << 8);
    idx = c;
    val = (float)s / (c == 0 ? 703.0f : 999.0f);
    return true;
  }
  if (c < 7) {
    const uint32_t t1 = a[11] & 15u, t2 = a[11] >> 4;
    idx = 5 + (int)(c == 5 ? t1 : t2);
    return c == 5 || t2 != t1;
  }
  if (c < 13) {
    const int k = c - 7;
    const uint32_t b = a[12] | (a[13] << 8) | (a[14] << 16);
    const int stage = boost_get(b, k);
    idx = 20 + k;
    val = ((float)boost_num(stage) / (float)boost_den(stage)) * (k < 4 ? 0.25f : (1.0f / 3.0f));
    return true;
  }
  uint64_t v;
  memcpy(&v, a + 16, 8);
  if (c < 32) {
    const int j = c - 13;  // 0..18
    idx = 26 + j;
    if (j < 16) {
      // bide thrashing charging binding invulnerable confusion mist focus_energy substitute recharging rage
      // leech_seed toxic light_screen reflect transform  (MultiHit = bit 2 and Flinch = bit 3 are not encoded)
      const int bit = j < 2 ? j : j + 2;
      return ((v >> bit) & 1ull) != 0;
    }
    if (j == 16) val = (float)vf_get(v, VF_STATE) / 65535.0f;
    else if (j == 17) val = (float)vf_get(v, VF_SUBSTITUTE) / 177.0f;
    else val = (float)vf_get(v, VF_TOXIC) / 16.0f;
    return val != 0.0f;
  }
  if (c < 36) {
    const uint32_t id = a[24 + 2 * (c - 32)], pp = a[24 + 2 * (c - 32) + 1];
    idx = 45 + (int)id - 1;
    return id != MOVE_None && id != MOVE_Struggle && pp != 0;
  }
  uint32_t n;
  if (c == 36) { n = bf_get(dur, DF_CONFUSION); idx = 209 + (int)n - 1; }
  else if (c == 37) { n = bf_get(dur, DF_DISABLE); idx = 214 + (int)n - 1; }
  else if (c == 38) { n = bf_get(dur, DF_ATTACKING); idx = 222 + (int)n - 1; }
  else { n = bf_get(dur, DF_BINDING); idx = 225 + (int)n - 1; }
  return n != 0;
}

// gather-sum first layer (affine.h:86-102): lane owns hidden units 4*lane..4*lane+3
__device__ __forceinline__ void sparse_layer(const float* __restrict__ WT, const float* __restrict__ bias, int H, const uint16_t* fi,
                                             const float* fv, int n, int act, float* h, int lane) {
  if (4 * lane < H) {
    float4 acc = *reinterpret_cast<const float4*>(bias + 4 * lane);
    for (int k = 0; k < n; k++) {
      const float v = fv[k];
      const float4 w = __ldg(reinterpret_cast<const float4*>(WT + (size_t)fi[k] * H + 4 * lane));
      acc.x = fmaf(v, w.x, acc.x);
      acc.y = fmaf(v, w.y, acc.y);
      acc.z = fmaf(v, w.z, acc.z);
      acc.w = fmaf(v, w.w, acc.w);
    }
    h[4 * lane + 0] = activate(acc.x, act);
    h[4 * lane + 1] = activate(acc.y, act);
    h[4 * lane + 2] = activate(acc.z, act);
    h[4 * lane + 3] = activate(acc.w, act);
  }
  __syncwarp();
}

// dense layer from a [K][O] matrix; lane owns outputs lane, lane+32, lane+64 (O <= 96)
__device__ __forceinline__ void dense_layer3(const float* WT, const float* bias, int K, int O, const float* x, int act, float* y, int lane) {
  float a0 = lane < O ? bias[lane] : 0.0f, a1 = lane + 32 < O ? bias[lane + 32] : 0.0f, a2 = lane + 64 < O ? bias[lane + 64] : 0.0f;
  for (int k = 0; k < K; k++) {
    const float xk = x[k];
    const float* row = WT + (size_t)k * O;
    if (lane < O) a0 = fmaf(xk, row[lane], a0);
    if (lane + 32 < O) a1 = fmaf(xk, row[lane + 32], a1);
    if (lane + 64 < O) a2 = fmaf(xk, row[lane + 64], a2);
  }
  if (lane < O) y[lane] = act ? activate(a0, act) : a0;
  if (lane + 32 < O) y[lane + 32] = act ? activate(a1, act) : a1;
  if (lane + 64 < O) y[lane + 64] = act ? activate(a2, act) : a2;
  __syncwarp();
}

__global__ void __launch_bounds__(EVAL_WARPS * 32)
eval_kernel(NetView nv, const uint8_t* __restrict__ battles, const uint8_t* __restrict__ durations, const uint8_t* __restrict__ results,
            float* __restrict__ values, float* __restrict__ logits, uint8_t* __restrict__ counts, uint32_t n) {
  extern __shared__ __align__(16) float smem[];
  const NetDims d = nv.d;
  // discrete mode: float embedding nets (clamp), embeddings and hp quantised to 0..127, integer main net (see activate)
  const bool disc = d.act == 3;
  const int act_h = disc ? 2 : d.act, act_e = d.act, act_m = disc ? 4 : d.act;
  const float out_div = disc ? 8128.0f : 1.0f;  // 127 * 64 (quantized/main-net.h:127)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // CTA-shared second-layer weights
  float* s_pk1T = smem;
  float* s_ac1T = s_pk1T + ((d.phd * d.pod + 3) & ~3);
  float* s_warp0 = s_ac1T + ((d.ahd * d.aod + 3) & ~3);
  for (int i = threadIdx.x; i < d.phd * d.pod; i += blockDim.x) s_pk1T[i] = nv.pk1T[i];
  for (int i = threadIdx.x; i < d.ahd * d.aod; i += blockDim.x) s_ac1T[i] = nv.ac1T[i];
  // per-warp scratch
  const int emb_pad = (d.in + 3) & ~3;
  const int per_warp = 96 + 4 + 128 + emb_pad + 64 * 5 + 64 + 32;  // floats
  float* w0 = s_warp0 + (size_t)warp * per_warp;
  uint32_t* rec = reinterpret_cast<uint32_t*>(w0);            // 96 words: the battle record
  uint32_t* durw = rec + 96;                                   // 2 words (+2 pad)
  float* h = w0 + 100;                                         // 128
  float* emb = h + 128;                                        // d.in
  float* b0 = emb + emb_pad;                                   // 64
  float* b1 = b0 + 64;                                         // 64
  float* hv = b1 + 64;                                         // 64 value hidden
  float* hp = hv + 64;                                         // 64 policy hidden (one side at a time)
  float* fv = hp + 64 + 64;                                    // 64 feature values
  uint16_t* fi = reinterpret_cast<uint16_t*>(fv + 64);         // 64 feature indices
  __syncthreads();

  const uint8_t* recb = reinterpret_cast<const uint8_t*>(rec);
  for (uint32_t e = blockIdx.x * EVAL_WARPS + warp; e < n; e += gridDim.x * EVAL_WARPS) {
    // ---- stage the record
    const uint32_t* src = reinterpret_cast<const uint32_t*>(battles + (size_t)e * 384);
    rec[lane] = src[lane];
    rec[lane + 32] = src[lane + 32];
    rec[lane + 64] = src[lane + 64];
    if (lane < 2) durw[lane] = reinterpret_cast<const uint32_t*>(durations + (size_t)e * 8)[lane];
    for (int i = lane; i < d.in; i += 32) emb[i] = 0.0f;
    __syncwarp();

    // ---- embeddings (network.h:131-175)
    for (int s = 0; s < 2; s++) {
      const uint8_t* side = recb + s * SIDE_SIZE;
      const uint32_t dur = durw[s];
      float* se = emb + s * d.side;
      for (int slot = 1; slot <= 6; slot++) {
        const uint32_t id = side[S_ORDER + slot - 1];
        if (id == 0) continue;
        const uint8_t* pk = side + (id - 1) * PK_SIZE;
        const uint32_t hpv = pk[PK_HP] | (pk[PK_HP + 1] << 8);
        if (hpv == 0) continue;
        const uint32_t maxhp = pk[0] | (pk[1] << 8);
        const uint32_t sleep = (dur >> (3 * (slot - 1))) & 7u;
        float* out = slot == 1 ? se : se + (1 + d.aod) + (slot - 2) * (1 + d.pod);
        // sparse feature list
        int cnt = 0;
        const int rounds = slot == 1 ? 2 : 1;
        for (int r = 0; r < rounds; r++) {
          const int c = lane + 32 * r;
          int idx = 0;
          float val = 0.0f;
          bool valid = false;
          if (slot == 1) {
            if (c < 40) valid = active_feature(side + S_ACTIVE, dur, c, idx, val);
            else if (c < 52) valid = pokemon_feature(pk, sleep, c - 40, 229, idx, val);
          } else if (c < 12) {
            valid = pokemon_feature(pk, sleep, c, 0, idx, val);
          }
          const uint32_t mask = __ballot_sync(0xFFFFFFFFu, valid);
          if (valid) {
            const int pos = cnt + __popc(mask & ((1u << lane) - 1u));
            fi[pos] = (uint16_t)idx;
            fv[pos] = val;
          }
          cnt += __popc(mask);
        }
        __syncwarp();
        if (slot == 1) {
          sparse_layer(nv.ac0T, nv.ac0b, d.ahd, fi, fv, cnt, act_h, h, lane);
          dense_layer3(s_ac1T, nv.ac1b, d.ahd, d.aod, h, act_e, out + 1, lane);
        } else {
          sparse_layer(nv.pk0T, nv.pk0b, d.phd, fi, fv, cnt, act_h, h, lane);
          dense_layer3(s_pk1T, nv.pk1b, d.phd, d.pod, h, act_e, out + 1, lane);
        }
        if (lane == 0) out[0] = disc ? floorf(((float)hpv / (float)maxhp) * 127.0f) : (float)hpv / (float)maxhp;
        __syncwarp();
      }
    }

    // ---- main net (main-net.h:66-107); lane owns outputs 2*lane, 2*lane+1
    {
      float a0 = 0.0f, a1 = 0.0f;
      const bool on = 2 * lane < d.hd;
      if (on) { a0 = nv.fc0b[2 * lane]; a1 = (2 * lane + 1 < d.hd) ? nv.fc0b[2 * lane + 1] : 0.0f; }
      if (on && !(d.hd & 1)) {
        for (int k = 0; k < d.in; k++) {
          const float x = emb[k];
          const float2 w = __ldg(reinterpret_cast<const float2*>(nv.fc0T + (size_t)k * d.hd + 2 * lane));
          a0 = fmaf(x, w.x, a0);
          a1 = fmaf(x, w.y, a1);
        }
      } else if (on) {
        for (int k = 0; k < d.in; k++) {
          const float x = emb[k];
          a0 = fmaf(x, nv.fc0T[(size_t)k * d.hd + 2 * lane], a0);
          if (2 * lane + 1 < d.hd) a1 = fmaf(x, nv.fc0T[(size_t)k * d.hd + 2 * lane + 1], a1);
        }
      }
      if (on) { b0[2 * lane] = activate(a0, act_m); if (2 * lane + 1 < d.hd) b0[2 * lane + 1] = activate(a1, act_m); }
      __syncwarp();
    }
    dense_layer3(nv.fc1T, nv.fc1b, d.hd, d.hd, b0, act_m, b1, lane);
    dense_layer3(nv.v2T, nv.v2b, d.hd, d.vhd, b1, act_m, hv, lane);
    {
      float part = 0.0f;
      for (int k = lane; k < d.vhd; k += 32) part = fmaf(hv[k], nv.v3w[k], part);
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
      if (lane == 0) values[e] = 1.0f / (1.0f + expf(-((part + nv.v3b[0]) / out_div)));
    }
    // ---- policy heads over the legal choices only
    Mem<1> m;
    m.set(const_cast<uint8_t*>(recb));
    const uint32_t res = results[e];
    for (int s = 0; s < 2; s++) {
      dense_layer3(s == 0 ? nv.p1aT : nv.p2aT, s == 0 ? nv.p1ab : nv.p2ab, d.hd, d.pohd, b1, act_m, hp, lane);
      const uint64_t list = mem_choices(m, s, (res >> (4 + 2 * s)) & 3);
      const int cnt = choice_count(list);
      const float* W = s == 0 ? nv.p1bW : nv.p2bW;
      const float* B = s == 0 ? nv.p1bb : nv.p2bb;
      const uint8_t* side = recb + s * SIDE_SIZE;
      for (int k = 0; k < 9; k++) {
        float lg = 0.0f;
        if (k < cnt) {
          const uint32_t c = choice_at(list, k);
          const uint32_t type = c & 3, data = c >> 2;
          uint32_t idx = 0;  // policy.h:29-58
          if (type == CH_MOVE) {
            // data == 0 (forced move): the reference reads moves[-1] = the low byte of stats.spc (policy.h:34-36), kept
            const uint8_t* stored = side + ((int)side[S_ORDER] - 1) * PK_SIZE;
            const uint32_t mid = data >= 1 ? stored[PK_MOVES + 2 * (data - 1)] : stored[PK_STATS + ST_SPC];
            idx = mid == 0 ? 0 : mid - 1;
          } else if (type == CH_SWITCH) {
            idx = 164 + side[((int)side[S_ORDER + data - 1] - 1) * PK_SIZE + PK_SPECIES] - 1;
          }
          float part = 0.0f;
          for (int j = lane; j < d.pohd; j += 32) part = fmaf(hp[j], W[(size_t)idx * d.pohd + j], part);
          for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
          lg = (part + B[idx]) / out_div;
        }
        if (lane == 0) logits[(size_t)e * 18 + s * 9 + k] = lg;
      }
      if (lane == 0) counts[(size_t)e * 2 + s] = (uint8_t)cnt;
      __syncwarp();
    }
  }
}

}  // namespace

#include "value_net_tc.cuh"

namespace {

uint16_t f2bf(float x) {  // round-to-nearest-even, as __float2bfloat16_rn
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)(u >> 16);
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
float bf2f(uint16_t h) {
  const uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
float rdf(const float* base, size_t i) {  // the parameter image need not be 4-byte aligned
  float v;
  memcpy(&v, reinterpret_cast<const uint8_t*>(base) + i * 4, 4);
  return v;
}

struct PackedMatrix {
  int N = 0;
  struct Piece { uint32_t gofs16, bytes16, ksteps, k0; };
  std::vector<Piece> pieces;
};

struct Packer {
  std::vector<uint16_t> w;
  std::vector<tc::Chunk> prog;

  // B[N][K] (K a multiple of 16, N of 16) -> split-bf16 chunks of at most one ring slot each
  template <class F>
  PackedMatrix matrix(int N, int K, F get) {
    PackedMatrix pm;
    pm.N = N;
    const int total = K / 16;
    int cs = (int)tc::RING_SLOT / (N * 64);
    if (cs < 1) cs = 1;
    for (int k0 = 0; k0 < total; k0 += cs) {
      const int ks = total - k0 < cs ? total - k0 : cs;
      const size_t base = w.size(), plane = (size_t)ks * N * 16;
      w.resize(base + 2 * plane);
      for (int kk = 0; kk < ks * 16; kk++)
        for (int nn = 0; nn < N; nn++) {
          const float v = get(nn, k0 * 16 + kk);
          const uint16_t hi = f2bf(v), lo = f2bf(v - bf2f(hi));
          const size_t idx = ((size_t)(kk >> 3) * N + nn) * 8 + (kk & 7);
          w[base + idx] = hi;
          w[base + plane + idx] = lo;
        }
      pm.pieces.push_back({(uint32_t)(base / 8), (uint32_t)(2 * plane / 8), (uint32_t)ks, (uint32_t)k0});
    }
    return pm;
  }

  // the same with the A operand in Tensor Memory: `operand` = tc::tmem_operand(first column, K steps per block), a_off16 =
  // the chunk's first K step; both planes exist for every step
  void stage_tmem(const PackedMatrix& pm, uint32_t operand, uint32_t dcol, bool overwrite, uint8_t end_flag, uint8_t begin_on) {
    for (size_t j = 0; j < pm.pieces.size(); j++) {
      const auto& pc = pm.pieces[j];
      tc::Chunk c;
      c.gofs16 = pc.gofs16;
      c.bytes16 = (uint16_t)pc.bytes16;
      c.ksteps = (uint8_t)pc.ksteps;
      c.n8 = (uint8_t)(pm.N / 8);
      c.a_off16 = (uint16_t)pc.k0;
      c.alo_off16 = (uint16_t)operand;
      c.lo_steps = (uint8_t)pc.ksteps;
      c.flags = (uint8_t)(tc::F_A_TMEM | begin_on | (j == 0 ? tc::F_STAGE_BEGIN : 0) | (j == 0 && overwrite ? tc::F_OVERWRITE : 0) |
                          (j + 1 == pm.pieces.size() ? end_flag : 0));
      c.dcol = (uint16_t)dcol;
      prog.push_back(c);
    }
  }

  void stage(const PackedMatrix& pm, uint32_t a_hi, uint32_t a_lo, int lo_ksteps, uint32_t dcol, bool overwrite, uint8_t end_flag,
             uint8_t begin_on = 0) {
    for (size_t j = 0; j < pm.pieces.size(); j++) {
      const auto& pc = pm.pieces[j];
      tc::Chunk c;
      c.gofs16 = pc.gofs16;
      c.bytes16 = (uint16_t)pc.bytes16;
      c.ksteps = (uint8_t)pc.ksteps;
      c.n8 = (uint8_t)(pm.N / 8);
      c.a_off16 = (uint16_t)((a_hi + pc.k0 * 2 * tc::CORE) / 16);
      c.alo_off16 = (uint16_t)((a_lo + pc.k0 * 2 * tc::CORE) / 16);
      int lo = lo_ksteps - (int)pc.k0;
      if (lo < 0) lo = 0;
      if (lo > (int)pc.ksteps) lo = (int)pc.ksteps;
      c.lo_steps = (uint8_t)lo;
      c.flags = (uint8_t)(begin_on | (j == 0 ? tc::F_STAGE_BEGIN : 0) | (j == 0 && overwrite ? tc::F_OVERWRITE : 0) |
                          (j + 1 == pm.pieces.size() ? end_flag : 0));
      c.dcol = (uint16_t)dcol;
      prog.push_back(c);
    }
  }
};

cudaError_t tc_pack(oakb200_net* n, const HostAffine* a, const NetDims& d) {
  n->tc_ok = d.phd <= 128 && d.ahd <= 128 && d.pod <= 63 && d.aod <= 95 && d.hd <= 64 && d.vhd <= 32 && d.pohd <= 64 && (d.pohd & 3) == 0;
  if (!n->tc_ok) return cudaSuccess;  // the fp32 kernel serves unusual shapes
  static_assert(sizeof(tc::Chunk) == 16, "chunk program entries are 16 bytes");
  Packer P;
  int inv_act[tc::K_ACT];
  for (int k = 0; k < tc::K_ACT; k++) inv_act[k] = -1;
  for (int idx = 0; idx < 229; idx++) inv_act[tc::act_col(idx)] = idx;
  const int phd = d.phd, pod = d.pod, ahd = d.ahd, aod = d.aod, hd = d.hd, in = d.in;
  const PackedMatrix ac0a = P.matrix(128, tc::K_ACT, [&](int o, int k) { return o < ahd && inv_act[k] >= 0 ? rdf(a[2].w, (size_t)o * 427 + inv_act[k]) : 0.0f; });
  int inv_pok[tc::K_POK];
  for (int k = 0; k < tc::K_POK; k++) inv_pok[k] = -1;
  for (int idx = 0; idx < 198; idx++) inv_pok[tc::pok_col(idx)] = idx;
  const PackedMatrix ac0p = P.matrix(128, tc::K_POK, [&](int o, int k) { return o < ahd && inv_pok[k] >= 0 ? rdf(a[2].w, (size_t)o * 427 + 229 + inv_pok[k]) : 0.0f; });
  const PackedMatrix pk0 = P.matrix(128, tc::K_POK, [&](int o, int k) { return o < phd && inv_pok[k] >= 0 ? rdf(a[0].w, (size_t)o * 198 + inv_pok[k]) : 0.0f; });
  const PackedMatrix pk1 = P.matrix(64, 128, [&](int o, int k) { return o < pod && k < phd ? rdf(a[1].w, (size_t)o * phd + k) : 0.0f; });
  const PackedMatrix ac1 = P.matrix(96, 128, [&](int o, int k) { return o < aod && k < ahd ? rdf(a[3].w, (size_t)o * ahd + k) : 0.0f; });
  // fc0 column slices: operand column k < od is embedding output k, column od is the hp fraction (network.h:131-175)
  auto fc0_slice = [&](int K, int od, int base) {
    return P.matrix(64, K, [&](int o, int k) {
      if (o >= hd || k > od) return 0.0f;
      return rdf(a[4].w, (size_t)o * in + (k < od ? base + 1 + k : base));
    });
  };
  PackedMatrix fa[2], fp[2][5];
  for (int s = 0; s < 2; s++) {
    fa[s] = fc0_slice(96, aod, s * d.side);
    for (int p = 0; p < 5; p++) fp[s][p] = fc0_slice(64, pod, s * d.side + (1 + aod) + p * (1 + pod));
  }
  const PackedMatrix fc1 = P.matrix(64, 64, [&](int o, int k) { return o < hd && k < hd ? rdf(a[5].w, (size_t)o * hd + k) : 0.0f; });
  const PackedMatrix heads = P.matrix(160, 64, [&](int o, int k) {
    if (k >= hd) return 0.0f;
    if (o < 32) return o < d.vhd ? rdf(a[6].w, (size_t)o * hd + k) : 0.0f;
    if (o < 96) return o - 32 < d.pohd ? rdf(a[8].w, (size_t)(o - 32) * hd + k) : 0.0f;
    return o - 96 < d.pohd ? rdf(a[10].w, (size_t)(o - 96) * hd + k) : 0.0f;
  });
  // The twelve party slots of a tile as one software pipeline (value_net_tc.cuh, worker loop): q = 6 s + slot - 1, slot 1 =
  // the active Pokemon. L1(0)a L1(0)b | L2(q) L1(q + 1) T(q) ...: every layer output is the next layer's A operand in place
  // in Tensor Memory; an active slot's first layer has two K parts (Active block, then its Pokemon block, accumulated).
  P.stage(ac0a, tc::A1_HI, tc::A1_LO, 1, tc::D_WORK, true, tc::F_END_D);
  P.stage(ac0p, tc::A1_HI, tc::A1_LO, 1, tc::D_WORK, false, tc::F_END_D);
  for (int q = 0; q < 12; q++) {
    const int s = q / 6, slot = q % 6;  // slot 0 = the active Pokemon
    if (slot == 0) P.stage_tmem(ac1, tc::tmem_operand(tc::D_WORK, tc::HID_NC / 16), tc::D_ACT, true, tc::F_END_E, tc::F_WAIT_A2);
    else P.stage_tmem(pk1, tc::tmem_operand(tc::D_WORK, tc::HID_NC / 16), tc::D_EMB, true, tc::F_END_E, tc::F_WAIT_A2);
    if (q == 5) P.stage(ac0a, tc::A1_HI, tc::A1_LO, 1, tc::D_WORK, true, tc::F_END_D);
    else if (q < 11) P.stage(pk0, tc::A1_HI, tc::A1_LO, 1, tc::D_WORK, true, tc::F_END_D);
    if (slot == 0) P.stage_tmem(fa[s], tc::tmem_operand(tc::D_ACT, tc::ACT_NC / 16), tc::D_TRUNK, s == 0, 0, tc::F_WAIT_E);
    else P.stage_tmem(fp[s][slot - 1], tc::tmem_operand(tc::D_EMB, tc::EMB_NC / 16), tc::D_TRUNK, false, q == 11 ? tc::F_END_T : 0, tc::F_WAIT_E);
    if (q == 5) P.stage(ac0p, tc::A1_HI, tc::A1_LO, 1, tc::D_WORK, false, tc::F_END_D);
  }
  P.stage_tmem(fc1, tc::tmem_operand(tc::D_TRUNK, tc::EMB_NC / 16), tc::D_WORK, true, tc::F_END_D, 0);
  P.stage_tmem(heads, tc::tmem_operand(tc::D_WORK, tc::EMB_NC / 16), tc::D_HEADS, true, tc::F_END_D, 0);
  if (getenv("OAKB200_EVAL_DEBUG"))
    for (size_t i = 0; i < P.prog.size(); i++)
      fprintf(stderr, "[prog %3zu] ksteps %u n %u a_off16 %u lo_steps %u flags 0x%02x dcol %u bytes %u\n", i, P.prog[i].ksteps,
              P.prog[i].n8 * 8, P.prog[i].a_off16, P.prog[i].lo_steps, P.prog[i].flags, P.prog[i].dcol, P.prog[i].bytes16 * 16);
  if (P.prog.size() > tc::PROG_MAX) {
    n->tc_ok = false;
    return cudaSuccess;
  }
  std::vector<float> c(tc::CONST_FLOATS, 0.0f);
  for (int i = 0; i < phd; i++) c[tc::C_PK0 + i] = rdf(a[0].bias, i);
  for (int i = 0; i < pod; i++) c[tc::C_PK1 + i] = rdf(a[1].bias, i);
  for (int i = 0; i < ahd; i++) c[tc::C_AC0 + i] = rdf(a[2].bias, i);
  for (int i = 0; i < aod; i++) c[tc::C_AC1 + i] = rdf(a[3].bias, i);
  for (int i = 0; i < hd; i++) c[tc::C_FC0 + i] = rdf(a[4].bias, i);
  for (int i = 0; i < hd; i++) c[tc::C_FC1 + i] = rdf(a[5].bias, i);
  for (int i = 0; i < d.vhd; i++) c[tc::C_HEADS + i] = rdf(a[6].bias, i);
  for (int i = 0; i < d.pohd; i++) c[tc::C_HEADS + 32 + i] = rdf(a[8].bias, i);
  for (int i = 0; i < d.pohd; i++) c[tc::C_HEADS + 96 + i] = rdf(a[10].bias, i);
  for (int i = 0; i < d.vhd; i++) c[tc::C_V3W + i] = rdf(a[7].w, i);
  c[tc::C_V3B] = rdf(a[7].bias, 0);
  cudaError_t e = cudaMalloc(&n->tc_wpack, P.w.size() * 2);
  if (e == cudaSuccess) e = cudaMemcpy(n->tc_wpack, P.w.data(), P.w.size() * 2, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&n->tc_prog, P.prog.size() * sizeof(tc::Chunk));
  if (e == cudaSuccess) e = cudaMemcpy(n->tc_prog, P.prog.data(), P.prog.size() * sizeof(tc::Chunk), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&n->tc_consts, c.size() * 4);
  if (e == cudaSuccess) e = cudaMemcpy(n->tc_consts, c.data(), c.size() * 4, cudaMemcpyHostToDevice);
  n->tc_nchunks = (int)P.prog.size();
  return e;
}

int launch_eval_tc(oakb200_ctx* ctx, size_t n, const void* battles, const void* durations, const void* results, void* value_out,
                   void* logits_out, void* counts_out) {
  const oakb200_net& N = *ctx->net;
  tc::Params P;
  P.wpack = N.tc_wpack;
  P.prog = reinterpret_cast<const tc::Chunk*>(N.tc_prog);
  P.consts = N.tc_consts;
  P.p1bW = N.blob + N.p1bW; P.p1bb = N.blob + N.p1bb; P.p2bW = N.blob + N.p2bW; P.p2bb = N.blob + N.p2bb;
  P.nchunks = N.tc_nchunks; P.act = N.d.act; P.pod = N.d.pod; P.aod = N.d.aod; P.vhd = N.d.vhd; P.pohd = N.d.pohd;
  const unsigned tiles = (unsigned)((n + tc::ROWS - 1) / tc::ROWS);
  // OAKB200_EVAL_CLUSTER=2: clusters of two CTAs share every weight chunk (one L2 read, multicast into both rings).
  // Bit-identical results and 20 % fewer L2 sector reads, but the same 0.535 ms per 64 Ki evaluations on B200
  // (profiles/r1l_eval_tc_cluster.txt): the weight stream is not what the tile's serial chain waits for. Not the default.
  static const int want_cluster = [] {
    const char* c = getenv("OAKB200_EVAL_CLUSTER");
    return c && c[0] == '2' ? 2 : 1;
  }();
  if (want_cluster == 2 && tiles >= 2) {
    auto k2 = tc::eval_tc_kernel<2>;
    CU(ctx, cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::SMEM_BYTES));
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(2, 1, 1);
    cfg.blockDim = dim3(tc::THREADS, 1, 1);
    cfg.dynamicSmemBytes = tc::SMEM_BYTES;
    cfg.stream = ctx->stream;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    static int max_clusters = -1;
    if (max_clusters < 0) {
      int mc = 0;
      if (cudaOccupancyMaxActiveClusters(&mc, k2, &cfg) != cudaSuccess) mc = 0;
      (void)cudaGetLastError();
      max_clusters = mc;
      if (getenv("OAKB200_EVAL_DEBUG")) fprintf(stderr, "[oakb200] eval_tc: %d resident clusters of 2\n", mc);
    }
    if (max_clusters > 0) {
      unsigned clusters = (tiles + 1) / 2;
      if (clusters > (unsigned)max_clusters) clusters = (unsigned)max_clusters;  // persistent: every cluster resident
      cfg.gridDim = dim3(2 * clusters, 1, 1);
      CU(ctx, cudaLaunchKernelEx(&cfg, k2, P, (const uint8_t*)battles, (const uint8_t*)durations, (const uint8_t*)results,
                                 (float*)value_out, (float*)logits_out, (uint8_t*)counts_out, (uint32_t)n));
      ctx->launches++;
      return OAKB200_OK;
    }
  }
  auto k1 = tc::eval_tc_kernel<1>;
  CU(ctx, cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::SMEM_BYTES));
  const unsigned grid = tiles < (unsigned)ctx->sm_count ? tiles : (unsigned)ctx->sm_count;  // persistent: one CTA per SM
  k1<<<grid, tc::THREADS, tc::SMEM_BYTES, ctx->stream>>>(P, (const uint8_t*)battles, (const uint8_t*)durations, (const uint8_t*)results,
                                                         (float*)value_out, (float*)logits_out, (uint8_t*)counts_out, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

}  // namespace

extern "C" int oakb200_eval_set_mode(oakb200_ctx* ctx, int mode) {
  if (!ctx || mode < 0 || mode > 2) return fail(ctx, OAKB200_EINVAL, "eval_set_mode: mode is 0 (auto), 1 (fp32 CUDA cores) or 2 (tensor cores)");
  ctx->eval_mode = mode;
  return OAKB200_OK;
}

extern "C" int oakb200_eval_device(oakb200_ctx* ctx, size_t n, const void* battles, const void* durations, const void* results,
                                   void* value_out, void* logits_out, void* counts_out) {
  if (!ctx || !battles || !durations || !results || !value_out || !logits_out || !counts_out)
    return fail(ctx, OAKB200_EINVAL, "eval: null argument");
  if (!ctx->net) return fail(ctx, OAKB200_EINVAL, "eval: no network loaded (oakb200_net_load)");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  const oakb200_net& N = *ctx->net;
  if (ctx->eval_mode == 2 && !N.tc_ok) return fail(ctx, OAKB200_EINVAL, "eval: network shape is outside the tensor-core kernel's limits");
  if (ctx->eval_mode != 1 && N.tc_ok) return launch_eval_tc(ctx, n, battles, durations, results, value_out, logits_out, counts_out);
  NetView v;
  v.d = N.d;
  const float* b = N.blob;
  v.pk0T = b + N.pk0T; v.pk0b = b + N.pk0b; v.pk1T = b + N.pk1T; v.pk1b = b + N.pk1b;
  v.ac0T = b + N.ac0T; v.ac0b = b + N.ac0b; v.ac1T = b + N.ac1T; v.ac1b = b + N.ac1b;
  v.fc0T = b + N.fc0T; v.fc0b = b + N.fc0b; v.fc1T = b + N.fc1T; v.fc1b = b + N.fc1b;
  v.v2T = b + N.v2T; v.v2b = b + N.v2b; v.v3w = b + N.v3w; v.v3b = b + N.v3b;
  v.p1aT = b + N.p1aT; v.p1ab = b + N.p1ab; v.p1bW = b + N.p1bW; v.p1bb = b + N.p1bb;
  v.p2aT = b + N.p2aT; v.p2ab = b + N.p2ab; v.p2bW = b + N.p2bW; v.p2bb = b + N.p2bb;
  const int emb_pad = (N.d.in + 3) & ~3;
  const size_t per_warp = 96 + 4 + 128 + emb_pad + 64 * 5 + 64 + 32;
  const size_t smem = (((size_t)N.d.phd * N.d.pod + 3) & ~(size_t)3) * 4 + (((size_t)N.d.ahd * N.d.aod + 3) & ~(size_t)3) * 4 +
                      per_warp * EVAL_WARPS * 4;
  CU(ctx, cudaFuncSetAttribute(eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, eval_kernel, EVAL_WARPS * 32, smem));
  if (per_sm < 1) per_sm = 1;
  unsigned grid = (unsigned)(ctx->sm_count * per_sm);
  const unsigned need = (unsigned)((n + EVAL_WARPS - 1) / EVAL_WARPS);
  if (grid > need) grid = need;
  eval_kernel<<<grid, EVAL_WARPS * 32, smem, ctx->stream>>>(v, (const uint8_t*)battles, (const uint8_t*)durations, (const uint8_t*)results,
                                                            (float*)value_out, (float*)logits_out, (uint8_t*)counts_out, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

extern "C" int oakb200_eval(oakb200_ctx* ctx, float* value_out, float* logits_out, uint8_t* counts_out) {
  if (!ctx) return OAKB200_EINVAL;
  const size_t n = ctx->n;
  int rc = oakb200_eval_device(ctx, n, ctx->battles, ctx->durations, ctx->results, ctx->eval_values, ctx->eval_logits, ctx->eval_counts);
  if (rc) return rc;
  if (value_out) CU(ctx, cudaMemcpyAsync(value_out, ctx->eval_values, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (logits_out) CU(ctx, cudaMemcpyAsync(logits_out, ctx->eval_logits, n * 72, cudaMemcpyDeviceToHost, ctx->stream));
  if (counts_out) CU(ctx, cudaMemcpyAsync(counts_out, ctx->eval_counts, n * 2, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}


// Leaf evaluation of a HOST batch with the transfers hidden: the batch is cut into chunks, every chunk's upload is queued
// on one copy stream, its evaluation on the context's stream behind that upload, and its download on a second copy
// stream behind the evaluation, so the two copy engines and the SMs work on three different chunks at a time. Same
// results as oakb200_upload + oakb200_eval (the kernels see the same records); the resident batch of the context is
// replaced by this one.
extern "C" int oakb200_eval_host(oakb200_ctx* ctx, size_t n, const uint8_t* battles, const uint8_t* durations, const uint8_t* results,
                                 float* value_out, float* logits_out, uint8_t* counts_out) {
  if (!ctx || !battles || !value_out || !logits_out || !counts_out) return fail(ctx, OAKB200_EINVAL, "eval_host: null argument");
  if (n > ctx->n_max) return fail(ctx, OAKB200_ECAPACITY, "eval_host: n exceeds n_max");
  if (!ctx->net) return fail(ctx, OAKB200_EINVAL, "eval: no network loaded (oakb200_net_load)");
  ctx->n = n;
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  const int MAX_CHUNKS = 64;
  if (!ctx->eval_pipe_ready) {
    CU(ctx, cudaStreamCreateWithFlags(&ctx->eval_h2d, cudaStreamNonBlocking));
    CU(ctx, cudaStreamCreateWithFlags(&ctx->eval_d2h, cudaStreamNonBlocking));
    for (int i = 0; i < 2 * MAX_CHUNKS + 1; i++) CU(ctx, cudaEventCreateWithFlags(&ctx->eval_ev[i], cudaEventDisableTiming));
    ctx->eval_pipe_ready = true;
  }
  size_t chunk = 148 * 128;  // one tile per SM of the tensor-core kernel
  if (const char* c = getenv("OAKB200_EVAL_CHUNK")) chunk = (size_t)atol(c) > 0 ? (size_t)atol(c) : chunk;  // tuning knob
  if ((n + chunk - 1) / chunk > (size_t)MAX_CHUNKS) chunk = (n + MAX_CHUNKS - 1) / MAX_CHUNKS;
  const int chunks = (int)((n + chunk - 1) / chunk);
  if (!durations) CU(ctx, cudaMemsetAsync(ctx->durations, 0, n * 8, ctx->stream));
  if (!results) CU(ctx, cudaMemsetAsync(ctx->results, 0x50, n, ctx->stream));
  CU(ctx, cudaEventRecord(ctx->eval_ev[0], ctx->stream));  // everything queued so far on the context's stream comes first
  CU(ctx, cudaStreamWaitEvent(ctx->eval_h2d, ctx->eval_ev[0], 0));
  CU(ctx, cudaStreamWaitEvent(ctx->eval_d2h, ctx->eval_ev[0], 0));
  for (int k = 0; k < chunks; k++) {
    const size_t at = (size_t)k * chunk, m = n - at < chunk ? n - at : chunk;
    CU(ctx, cudaMemcpyAsync(ctx->battles + at * 384, battles + at * 384, m * 384, cudaMemcpyHostToDevice, ctx->eval_h2d));
    if (durations) CU(ctx, cudaMemcpyAsync(ctx->durations + at * 8, durations + at * 8, m * 8, cudaMemcpyHostToDevice, ctx->eval_h2d));
    if (results) CU(ctx, cudaMemcpyAsync(ctx->results + at, results + at, m, cudaMemcpyHostToDevice, ctx->eval_h2d));
    CU(ctx, cudaEventRecord(ctx->eval_ev[1 + 2 * k], ctx->eval_h2d));
  }
  for (int k = 0; k < chunks; k++) {
    const size_t at = (size_t)k * chunk, m = n - at < chunk ? n - at : chunk;
    CU(ctx, cudaStreamWaitEvent(ctx->stream, ctx->eval_ev[1 + 2 * k], 0));
    const int rc = oakb200_eval_device(ctx, m, ctx->battles + at * 384, ctx->durations + at * 8, ctx->results + at, ctx->eval_values + at,
                                       ctx->eval_logits + at * 18, ctx->eval_counts + at * 2);
    if (rc) return rc;
    CU(ctx, cudaEventRecord(ctx->eval_ev[2 + 2 * k], ctx->stream));
    CU(ctx, cudaStreamWaitEvent(ctx->eval_d2h, ctx->eval_ev[2 + 2 * k], 0));
    CU(ctx, cudaMemcpyAsync(value_out + at, ctx->eval_values + at, m * 4, cudaMemcpyDeviceToHost, ctx->eval_d2h));
    CU(ctx, cudaMemcpyAsync(logits_out + at * 18, ctx->eval_logits + at * 18, m * 18 * 4, cudaMemcpyDeviceToHost, ctx->eval_d2h));
    CU(ctx, cudaMemcpyAsync(counts_out + at * 2, ctx->eval_counts + at * 2, m * 2, cudaMemcpyDeviceToHost, ctx->eval_d2h));
  }
  CU(ctx, cudaStreamSynchronize(ctx->eval_d2h));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}
