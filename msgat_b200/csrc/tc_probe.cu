// tc_probe.cu -- self-test of the tcgen05 building block (tests/test_gpu_tc.py): one CTA stages
// two SW128 tiles, issues kind::tf32 MMAs into TMEM and reads the accumulator back.
//   mode 0 (MN-major operands, the channel mixes):   D[p, o] = sum_k A[k, p] * B[k, o]
//          A [K][256] positions contiguous, B [K][N]; D [256][N]   (two M = 128 blocks)
//   mode 1 (K-major operands, the weight gradients): D[m, n] = sum_p A[m, p] * B[n, p]
//          A [Ra][P], B [Rb][P], P = 32 * npb positions; D [128][N] (rows >= Ra / cols >= Rb are junk)
#include "kernels.h"
#include "tc_common.cuh"

struct ProbeArgs {
  int mode, K, N, Ra, Rb, P;
  const float *A, *B;
  float* D;
};

__global__ void __launch_bounds__(128, 1) k_tc_probe(ProbeArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t sA = smem_u32(smem);
  // zero everything the MMAs may touch so that junk rows are finite
  for (int i = tid; i < 200 * 1024 / 16; i += blockDim.x) reinterpret_cast<float4*>(smem)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  // layouts: 0 = SW128 (16 B chunks), 2 = unswizzled core matrices, 3 = SW128 with 32 B atoms
  const int lay = (a.mode == 0 || a.mode == 1) ? 0 : (a.mode == 3 ? 3 : 2);
  const bool mn = (a.mode == 0 || a.mode == 2 || a.mode == 3);
  const int RA = mn ? a.K : a.Ra, RB = mn ? a.K : a.Rb, WA = mn ? 256 : a.P, WB = mn ? a.N : a.P;
  uint32_t cbA, cbB;
  const uint32_t offB = 96u * 1024u;
  if (lay == 2) { cbA = (uint32_t)WA * 32u; cbB = (uint32_t)WB * 32u; }          // row-group stride
  else { cbA = (uint32_t)((RA + 7) / 8) * 1024u; cbB = (uint32_t)((RB + 7) / 8) * 1024u; }  // column-block stride
  for (int i = tid; i < RA * WA; i += blockDim.x) {
    const int r = i / WA, p = i - r * WA;
    const uint32_t o = lay == 0 ? sw128_off(r, p, cbA) : lay == 3 ? sw128b32_off(r, p, cbA) : core_off(r, p, cbA);
    *reinterpret_cast<float*>(smem + o) = a.A[i];
  }
  for (int i = tid; i < RB * WB; i += blockDim.x) {
    const int r = i / WB, p = i - r * WB;
    const uint32_t o = lay == 0 ? sw128_off(r, p, cbB) : lay == 3 ? sw128b32_off(r, p, cbB) : core_off(r, p, cbB);
    *reinterpret_cast<float*>(smem + offB + o) = a.B[i];
  }
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_slot), 256);
  fence_smem_to_async();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = tmem_slot;
  if (tid == 0) {
    if (mn) {
      const uint32_t idesc = umma_idesc_tf32(128, a.N, 1, 1);
      for (int mb = 0; mb < 2; ++mb)
        for (int ks = 0; ks < a.K / 8; ++ks) {
          uint64_t da, db;
          if (lay == 2) {  // MN-major, no swizzle: SBO = next 4 positions, LBO = next 8 rows
            da = umma_desc(sA + (uint32_t)mb * 32u * 128u + (uint32_t)ks * cbA, cbA, 128u, 0);
            db = umma_desc(sA + offB + (uint32_t)ks * cbB, cbB, 128u, 0);
          } else if (lay == 3) {  // 4-row atoms: one k-step = 2 atoms (SBO = 512), column blocks LBO apart
            da = umma_desc(sA + (uint32_t)mb * 4u * cbA + (uint32_t)ks * 1024u, cbA, 512u, 1);
            db = umma_desc(sA + offB + (uint32_t)ks * 1024u, cbB, 512u, 1);
          } else {
            da = umma_desc(sA + (uint32_t)mb * 4u * cbA + (uint32_t)ks * 1024u, cbA, 1024u, 2);
            db = umma_desc(sA + offB + (uint32_t)ks * 1024u, cbB, 1024u, 2);
          }
          umma_tf32(tm + (uint32_t)mb * 128u, da, db, idesc, ks > 0);
        }
    } else {
      const uint32_t idesc = umma_idesc_tf32(128, a.N, 0, 0);
      int first = 1;
      for (int k8 = 0; k8 < a.P / 8; ++k8) {
        uint64_t da, db;
        if (lay == 2) {  // K-major, no swizzle: SBO = next 8 rows, LBO = next 4 positions
          da = umma_desc(sA + (uint32_t)k8 * 256u, 128u, cbA, 0);
          db = umma_desc(sA + offB + (uint32_t)k8 * 256u, 128u, cbB, 0);
        } else {
          da = umma_desc(sA + (uint32_t)(k8 >> 2) * cbA + (uint32_t)(k8 & 3) * 32u, 0u, 1024u, 2);
          db = umma_desc(sA + offB + (uint32_t)(k8 >> 2) * cbB + (uint32_t)(k8 & 3) * 32u, 0u, 1024u, 2);
        }
        umma_tf32(tm, da, db, idesc, first ? 0u : 1u);
        first = 0;
      }
    }
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after_sync();
  const int nblk = mn ? 2 : 1;
  for (int mb = 0; mb < nblk; ++mb)
    for (int c0 = 0; c0 < a.N; c0 += 8) {
      float v[8];
      tmem_ld8(tmem_addr(tm, warp * 32, mb * 128 + c0), v);
      for (int j = 0; j < 8; ++j) a.D[((int64_t)mb * 128 + tid) * a.N + c0 + j] = v[j];
    }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 256);
}

extern "C" int msgat_tc_probe(int mode, const float* A, const float* B, float* D, int K, int N, int Ra, int Rb, int P,
                              void* stream) {
  const bool mnm = mode == 0 || mode == 2 || mode == 3;
  if (mnm && (K % 8 || K > 72 || N % 16 || N > 96)) { msgat_set_error("tc_probe: bad shape"); return -1; }
  if (!mnm && (P % 32 || P > 192 || Ra > 96 || Rb > 96 || N % 16 || N > 128)) { msgat_set_error("tc_probe: bad shape"); return -1; }
  ProbeArgs a{mode, K, N, Ra, Rb, P, A, B, D};
  cudaFuncSetAttribute(k_tc_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  k_tc_probe<<<1, 128, 200 * 1024, (cudaStream_t)stream>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { msgat_set_error("tc_probe: %s", cudaGetErrorString(e)); return -100; }
  return 0;
}
