/* dartray_hash.h — order-independent checksum of an intersection list.
 * Shared by the CUDA engine and the CPU oracle (test infrastructure) so that full-size
 * runs can be compared without storing the list: the checksum is the wrapping 64-bit sum
 * of dre_isect_mix() over every ray-cell intersection (one iteration of the main loop at
 * rt_routines.f90:1689). */
#ifndef DARTRAY_HASH_H
#define DARTRAY_HASH_H
#include <stdint.h>

#if defined(__CUDACC__)
#define DRE_HD __host__ __device__ inline
#else
#define DRE_HD static inline
#endif

DRE_HD uint64_t dre_mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

DRE_HD uint64_t dre_isect_mix(int32_t unit, int32_t step, int32_t ray, int32_t k,
                              int32_t cell, int32_t isel, uint64_t psel_bits) {
  uint64_t h = dre_mix64(((uint64_t)(uint32_t)unit << 32) | (uint32_t)step);
  h = dre_mix64(h ^ (((uint64_t)(uint32_t)ray << 32) | (uint32_t)k));
  h = dre_mix64(h ^ (((uint64_t)(uint32_t)cell << 8) | (uint32_t)isel));
  h = dre_mix64(h ^ psel_bits);
  return h;
}
#endif
