// Operand images of the tensor-core path (shared by contract.cu, update.cu, tc_path.cu).
//
// Both images are the exact byte order the UMMA shared-memory descriptor expects for the
// SWIZZLE_NONE ("interleave") canonical layouts, so a K-chunk is one contiguous block that a
// single 1-D bulk copy (cp.async.bulk) drops into shared memory — no tensor map needed.
// A core matrix is 8 x 16 bytes = 128 contiguous bytes (8 fp16 per 16-byte row).
//
//   Uimg  (forward B operand, K-major):  N index nn in [0, 2*rpad): nn <  rpad -> hi(U'[j, nn])
//                                                                 nn >= rpad -> lo(U'[j, nn-rpad])
//          per k:  [j/8][nn/8][nn%8][j%8]      K = feature j (npad, multiple of 64), U' = U * 2^uexp
//   Eimg  (backward B operand, MN-major): same nn convention for E'[i, l] = E[k][l][i] * escale
//          per k:  [i/8][nn/8][i%8][nn%8]      K = minibatch row i (mpad, multiple of 128)
//
// Descriptor strides for both: SBO (next core matrix along N) = 128 B,
//                              LBO (next core matrix along K) = (2*rpad/8) * 128 B.
#pragma once
#include <stdint.h>

__host__ __device__ inline size_t gptb_uimg_offset(int k, int j, int nn, int npad, int rpad) {
    return (size_t)k * npad * 2 * rpad + ((size_t)(j >> 3) * (2 * rpad / 8) + (nn >> 3)) * 64 + (nn & 7) * 8 + (j & 7);
}
// offset of element (row i, nn = 0); core matrix bq along N adds bq*64, element l%8 adds itself.
__host__ __device__ inline size_t gptb_eimg_offset(int k, int64_t i, int64_t mpad, int rpad) {
    return (size_t)k * mpad * 2 * rpad + (size_t)(i >> 3) * (2 * rpad / 8) * 64 + (size_t)(i & 7) * 8;
}
