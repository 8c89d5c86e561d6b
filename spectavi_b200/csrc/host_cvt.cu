// host_cvt.cu -- the AVX2 body of the host-side float32 -> uint8 conversion (host_ctx.cu f32_to_u8_host).  This file,
// and only this file, is compiled with -mavx2 (Makefile); the caller checks __builtin_cpu_supports("avx2") first.
// Semantics = K0 (cascade.cu f32_to_u8_kernel, CascadingHashNn.h:236-239): (uint8)(cvt.rzi.s32.f32(v) + 128), where the
// device conversion saturates and maps NaN to 0.
#include <immintrin.h>

#include <cstddef>
#include <cstdint>

namespace spv {

// n % 32 == 0
void f32_to_u8_avx2(const float* s, uint8_t* d, size_t n) {
  const __m256i k128 = _mm256_set1_epi32(128), kff = _mm256_set1_epi32(255), perm = _mm256_setr_epi32(0, 4, 1, 5, 2, 6, 3, 7);
  const __m256 big = _mm256_set1_ps(2147483648.0f);
  // cvttps gives 0x80000000 for NaN and for anything outside int32: INT_MIN, which is what the device gives for
  // v <= -2^31; NaN (device: 0) has the same low byte after + 128; v >= 2^31 (device: INT_MAX) is fixed by the XOR
  auto one = [&](const float* p) {
    const __m256 v = _mm256_loadu_ps(p);
    __m256i r = _mm256_cvttps_epi32(v);
    r = _mm256_xor_si256(r, _mm256_castps_si256(_mm256_cmp_ps(v, big, _CMP_GE_OQ)));
    return _mm256_and_si256(_mm256_add_epi32(r, k128), kff);
  };
  for (size_t i = 0; i < n; i += 32) {
    const __m256i q = _mm256_packus_epi16(_mm256_packus_epi32(one(s + i), one(s + i + 8)),
                                          _mm256_packus_epi32(one(s + i + 16), one(s + i + 24)));
    _mm256_storeu_si256(reinterpret_cast<__m256i*>(d + i), _mm256_permutevar8x32_epi32(q, perm));
  }
}

}  // namespace spv
