// Epilogue of the token-major tcgen05 GEMMs (gemm_big.cu, gemm_pair.cu): 32 consecutive output features of one token,
// accumulator bits in r[] -> bias / GELU(erf) / bf16 round / LayerScale / residual -> fp32, fp16 or bf16 store.
#pragma once
#include "gemm.cuh"
#include "ptx.cuh"

namespace mt {

// GELU(erf) of the token-major epilogues.  erff() costs ~40 instructions with two divergent branches, and the codec's
// linear1 epilogue (128 x 256 erf evaluations per tile on 8 warps) was bound by it (+150 us on a 250 us GEMM).  Branch-free
// form: erf(|z|) = 1 - 2^(-q(|z|)), q = -log2(erfc) fitted by a degree-8 polynomial with q(0) = 0 on [0, 4] (erf(4) = 1 in
// fp32), weighted so that the ERROR OF erf is uniform: 1.6e-8 in exact arithmetic, 1.2e-7 evaluated in fp32.  GELU max abs
// error over [-10, 10] vs float64: 4.7e-7 (torch's own fp32 GELU: 1.2e-6).  16 instructions.
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float z = x * 0.70710678118654752440f;
  const float t = fminf(fabsf(z), 4.0f);
  float q = 4.535859748e-05f;
  q = fmaf(q, t, -4.455076315e-04f);
  q = fmaf(q, t, 1.489441567e-03f);
  q = fmaf(q, t, 7.746297760e-04f);
  q = fmaf(q, t, -2.825368119e-02f);
  q = fmaf(q, t, 1.484816151e-01f);
  q = fmaf(q, t, 9.184163932e-01f);
  q = fmaf(q, t, 1.627908593e+00f);
  q *= -t;
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(q));
  return 0.5f * x * (1.f + copysignf(1.f - e, z));
}

__device__ __forceinline__ void epi_chunk32(const Epi& ep, int m, int n0, int N, const uint32_t (&r)[32]) {
  const bool full_chunk = n0 + 32 <= N;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  if (ep.bias) {
#pragma unroll
    for (int j = 0; j < 32; ++j) if (full_chunk || n0 + j < N) v[j] += ep.bias[n0 + j];
  }
  if (ep.act == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
  }
  if (ep.round_bf16) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = rbf(v[j]);
  }
  if (ep.scale) {
#pragma unroll
    for (int j = 0; j < 32; ++j) if (full_chunk || n0 + j < N) v[j] *= ep.scale[n0 + j];
  }
  if (ep.residual) {
    if (ep.out_f32 == 1) {
      const float* rr = reinterpret_cast<const float*>(ep.residual) + static_cast<long long>(m) * ep.ldr + n0;
      if (full_chunk && (ep.ldr & 3) == 0) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 x = *reinterpret_cast<const float4*>(rr + j);
          v[j] = x.x + v[j]; v[j + 1] = x.y + v[j + 1]; v[j + 2] = x.z + v[j + 2]; v[j + 3] = x.w + v[j + 3];
        }
      } else {
        for (int j = 0; j < 32; ++j) if (n0 + j < N) v[j] = rr[j] + v[j];
      }
    } else if (ep.out_f32 == 2) {
      const __half* rr = reinterpret_cast<const __half*>(ep.residual) + static_cast<long long>(m) * ep.ldr + n0;
      for (int j = 0; j < 32; ++j) if (full_chunk || n0 + j < N) v[j] = __half2float(rr[j]) + v[j];
    } else {
      const bf16* rr = reinterpret_cast<const bf16*>(ep.residual) + static_cast<long long>(m) * ep.ldr + n0;
      for (int j = 0; j < 32; ++j) if (full_chunk || n0 + j < N) v[j] = bf2f(rr[j]) + v[j];
    }
  }
  if (ep.out_f32 == 1) {
    float* dst = reinterpret_cast<float*>(ep.out) + static_cast<long long>(m) * ep.ldo + n0;
    if (full_chunk && (ep.ldo & 3) == 0) {
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
      for (int j = 0; j < 32; ++j) if (n0 + j < N) dst[j] = v[j];
    }
  } else if (ep.out_f32 == 2) {
    __half* dst = reinterpret_cast<__half*>(ep.out) + static_cast<long long>(m) * ep.ldo + n0;
    if (full_chunk && (ep.ldo & 15) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 31) == 0) {
      // 32-byte stores (sm_100): every store writes a whole sector of this lane's row
#pragma unroll
      for (int j = 0; j < 32; j += 16) {
        uint32_t pk[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) pk[e] = pack_f16(v[j + 2 * e], v[j + 2 * e + 1]);
        asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + j), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]),
                     "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
                     : "memory");
      }
    } else if (full_chunk && (ep.ldo & 7) == 0) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 pk;
        pk.x = pack_f16(v[j], v[j + 1]); pk.y = pack_f16(v[j + 2], v[j + 3]);
        pk.z = pack_f16(v[j + 4], v[j + 5]); pk.w = pack_f16(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4*>(dst + j) = pk;
      }
    } else {
      for (int j = 0; j < 32; ++j) if (n0 + j < N) dst[j] = __float2half_rn(v[j]);
    }
  } else {
    bf16* dst = reinterpret_cast<bf16*>(ep.out) + static_cast<long long>(m) * ep.ldo + n0;
    if (full_chunk && (ep.ldo & 7) == 0) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 pk;
        pk.x = pack_bf16(v[j], v[j + 1]); pk.y = pack_bf16(v[j + 2], v[j + 3]);
        pk.z = pack_bf16(v[j + 4], v[j + 5]); pk.w = pack_bf16(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4*>(dst + j) = pk;
      }
    } else {
      for (int j = 0; j < 32; ++j) if (n0 + j < N) dst[j] = f2bf(v[j]);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Coalesced form of the same chunk for a whole warp.  tcgen05.ld hands every lane ONE token row (32 consecutive features),
// so the direct form above touches 32 different rows with every 16-byte load / store.  Here the warp transposes its
// 32 x 32 chunk through a swizzled shared-memory tile and then walks it in row-major order: every global access of the
// warp covers whole 128-byte (fp32) / 64-byte (16-bit) row segments, the residual is read the same way, and its loads are
// issued a chunk ahead (EpiRes) so their latency is hidden.  Arithmetic per element is the direct form's, in the same
// order (bit-identical results).  Needs n0 + 32 <= N and 16-byte aligned rows (epi_coalesced_ok).
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kEpiTileFloats = 32 * 32;       // 32 rows x 8 sixteen-byte chunks, chunk index XOR (row & 7): conflict-free both ways

__device__ __forceinline__ bool epi_coalesced_ok(const Epi& ep) {
  const int per16 = ep.out_f32 == 1 ? 4 : 8;       // elements per 16 bytes
  // GELU (the codec's linear1) is bound by the erf arithmetic of its 8 epilogue warps, not by its stores: measured 15 % slower
  // through the transposition (profiles/r02_gemm_pair_ab.jsonl), so it keeps the per-lane form
  bool ok = !ep.direct_store && ep.act != 1 && (ep.ldo % per16) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0;
  if (ep.residual) ok = ok && (ep.ldr % per16) == 0 && (reinterpret_cast<uintptr_t>(ep.residual) & 15) == 0;
  return ok;
}

// residual of one 32 x 32 chunk, row-major over the warp.  The kernels request chunk c + 1 (and chunk 0 of the next tile,
// before they wait for its accumulator) while chunk c is processed: the loads have a whole chunk of work to land behind.
struct EpiRes {
  uint4 w[8];        // fp32 residual: 8 x float4 (rows i * 4 + lane / 8); 16-bit residual: w[0..3] (rows i * 8 + lane / 4)
};
__device__ __forceinline__ void epi_res_load(const Epi& ep, int m0, int n0, int M, int lane, EpiRes& res) {
  if (ep.rope) {
    // rotary epilogue (16-bit output, no residual): the (cos, sin) pairs of the lane's four (row, 8-column group) cells
    if (n0 < ep.rope_cols) {
      const int grp = ((n0 + (lane & 3) * 8) & 63) >> 3;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = m0 + i * 8 + (lane >> 2);
        if (m < M) {
          const float4* cs = ep.rope + static_cast<long long>(m % ep.rope_T) * 16 + 2 * grp;
          res.w[2 * i] = *reinterpret_cast<const uint4*>(cs);
          res.w[2 * i + 1] = *reinterpret_cast<const uint4*>(cs + 1);
        }
      }
    }
    return;
  }
  if (!ep.residual) return;
  if (ep.out_f32 == 1) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int m = m0 + i * 4 + (lane >> 3);
      if (m < M)
        res.w[i] = *reinterpret_cast<const uint4*>(reinterpret_cast<const float*>(ep.residual) + static_cast<long long>(m) * ep.ldr + n0 + (lane & 7) * 4);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + i * 8 + (lane >> 2);
      if (m < M)
        res.w[i] = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(ep.residual) + static_cast<long long>(m) * ep.ldr + n0 + (lane & 3) * 8);
    }
  }
}

__device__ __forceinline__ void epi_chunk32_warp(const Epi& ep, int m0, int n0, int M, const uint32_t (&r)[32], float* tile,
                                                 int lane, const EpiRes& res) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  if (ep.bias) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] += ep.bias[n0 + j];
  }
  if (ep.act == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
  }
  if (ep.round_bf16) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = rbf(v[j]);
  }
  if (ep.scale) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= ep.scale[n0 + j];
  }
  // lane = token row -> shared tile (16-byte chunk c of row r lives at chunk c ^ (r & 7))
  float4* tile4 = reinterpret_cast<float4*>(tile);
#pragma unroll
  for (int c = 0; c < 8; ++c) tile4[lane * 8 + (c ^ (lane & 7))] = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
  __syncwarp();
  if (ep.out_f32 == 1) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int row = i * 4 + (lane >> 3), c = (lane & 7) * 4;
      const int m = m0 + row;
      float4 a = tile4[row * 8 + ((lane & 7) ^ (row & 7))];
      if (m < M) {
        if (ep.residual) {
          a.x = __uint_as_float(res.w[i].x) + a.x; a.y = __uint_as_float(res.w[i].y) + a.y;
          a.z = __uint_as_float(res.w[i].z) + a.z; a.w = __uint_as_float(res.w[i].w) + a.w;
        }
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(ep.out) + static_cast<long long>(m) * ep.ldo + n0 + c) = a;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int row = i * 8 + (lane >> 2), c = (lane & 3) * 8;
      const int m = m0 + row;
      const float4 lo = tile4[row * 8 + ((2 * (lane & 3)) ^ (row & 7))], hi = tile4[row * 8 + ((2 * (lane & 3) + 1) ^ (row & 7))];
      float a[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
      if (m < M) {
        if (ep.rope && n0 + c < ep.rope_cols) {
          // rotary embedding of a q | k column group (interleaved pairs, 64-wide heads): the in_proj epilogue of the codec
          // rotates the fp32 accumulators, one rounding to half instead of the separate pass's two
          const float4 cs0 = *reinterpret_cast<const float4*>(&res.w[2 * i]), cs1 = *reinterpret_cast<const float4*>(&res.w[2 * i + 1]);
          const float x0 = a[0], x1 = a[1], x2 = a[2], x3 = a[3], x4 = a[4], x5 = a[5], x6 = a[6], x7 = a[7];
          a[0] = x0 * cs0.x - x1 * cs0.y; a[1] = x0 * cs0.y + x1 * cs0.x;
          a[2] = x2 * cs0.z - x3 * cs0.w; a[3] = x2 * cs0.w + x3 * cs0.z;
          a[4] = x4 * cs1.x - x5 * cs1.y; a[5] = x4 * cs1.y + x5 * cs1.x;
          a[6] = x6 * cs1.z - x7 * cs1.w; a[7] = x6 * cs1.w + x7 * cs1.z;
        }
        if (ep.residual && !ep.rope) {
          const uint32_t w[4] = {res.w[i].x, res.w[i].y, res.w[i].z, res.w[i].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (ep.out_f32 == 2) {
              const __half2 h = *reinterpret_cast<const __half2*>(&w[e]);
              a[2 * e] = __low2float(h) + a[2 * e]; a[2 * e + 1] = __high2float(h) + a[2 * e + 1];
            } else {
              a[2 * e] = bits_lo(w[e]) + a[2 * e]; a[2 * e + 1] = bits_hi(w[e]) + a[2 * e + 1];
            }
          }
        }
        uint4 pk;
        if (ep.out_f32 == 2) { pk.x = pack_f16(a[0], a[1]); pk.y = pack_f16(a[2], a[3]); pk.z = pack_f16(a[4], a[5]); pk.w = pack_f16(a[6], a[7]); }
        else { pk.x = pack_bf16(a[0], a[1]); pk.y = pack_bf16(a[2], a[3]); pk.z = pack_bf16(a[4], a[5]); pk.w = pack_bf16(a[6], a[7]); }
        *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(ep.out) + static_cast<long long>(m) * ep.ldo + n0 + c) = pk;
      }
    }
  }
  __syncwarp();                                 // the tile is rewritten by the next chunk
}

// One accumulator tile of an epilogue warp: its 32 token rows (m0 ..) x COLS features from n_base, accumulator at `taddr`.
// `cur` holds the residual of chunk 0, requested by the caller BEFORE it waited for the accumulator.
template <int COLS>
__device__ __forceinline__ void epi_tile_warp(const Epi& ep, bool coalesced, int m0, int n_base, int M, int N, uint32_t taddr,
                                              float* tile, int lane, EpiRes& cur) {
#pragma unroll 1
  for (int c0 = 0; c0 < COLS; c0 += 32) {
    const int n0 = n_base + c0;
    EpiRes nxt = cur;
    if (c0 + 32 < COLS && coalesced && n0 + 64 <= N && m0 < M) epi_res_load(ep, m0, n0 + 32, M, lane, nxt);
    uint32_t r[32];
    ptx::tmem_ld32(taddr + c0, r);
    ptx::tmem_ld_wait();
    if (coalesced && n0 + 32 <= N) { if (m0 < M) epi_chunk32_warp(ep, m0, n0, M, r, tile, lane, cur); }
    else if (m0 + lane < M && n0 < N) epi_chunk32(ep, m0 + lane, n0, N, r);
    cur = nxt;
  }
}

}  // namespace mt
