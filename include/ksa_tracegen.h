/*
 * ksa_tracegen.h — the synthetic queue-depth generator, as integer-only inline functions that
 * compile unchanged as C (oracle/, host tools) and as CUDA device code (csrc/ksa_gen.cu), so the
 * CPU baseline and the GPU replay see the same bits for the same (seed, trace, tick).
 *
 * This is input synthesis, not part of the reference: the reference's only depth source is the
 * SQS API (sqs/sqs.go:51) or the constant MockSQS fixture of its tests (main_test.go:288-304).
 * Ticks are absolute and 1-based (tick k is the k-th poll, main.go:41); trace ids are global, so
 * sharding traces over GPUs or chunking ticks never changes a value.
 */
#ifndef KSA_TRACEGEN_H
#define KSA_TRACEGEN_H

#include <stdint.h>
#include "ksa_replay.h"

#if defined(__CUDACC__)
#define KSA_HD __host__ __device__ __forceinline__
#else
#define KSA_HD static inline
#endif

/* splitmix64 finaliser (Steele, Lea, Flood 2014): a bijection on 64-bit words. */
KSA_HD uint64_t ksa_mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

/* Per-(seed, attribute-slot, trace) key; one more mix with the tick gives the per-entry word. */
KSA_HD uint64_t ksa_trace_key(uint64_t seed, uint32_t slot, uint64_t trace) {
  return ksa_mix64(ksa_mix64(seed ^ ((uint64_t)slot << 56)) + trace);
}
KSA_HD uint64_t ksa_entry_word(uint64_t key, uint64_t tick) {
  return ksa_mix64(key + tick * 0xD1342543DE82EF95ull);
}

/* Uniform integer in [0, range) from the top 32 bits of h (range <= 2^32). */
KSA_HD uint64_t ksa_below(uint64_t h, uint64_t range) { return ((h >> 32) * range) >> 32; }

/* BANDS: 40 % in [0, low_max], 30 % in (low_max, high_min), 30 % in [high_min, high_max]. */
KSA_HD int64_t ksa_gen_bands_value(const ksa_gen_params* g, uint64_t h) {
  uint32_t band = (uint32_t)(((h & 0xFFFFu) * 10u) >> 16); /* 0..9 */
  if (band < 4) return (int64_t)ksa_below(h, (uint64_t)g->low_max + 1);
  if (band < 7) {
    uint64_t width = (uint64_t)(g->high_min - g->low_max - 1);
    if (width == 0) return g->low_max;
    return g->low_max + 1 + (int64_t)ksa_below(h, width);
  }
  return g->high_min + (int64_t)ksa_below(h, (uint64_t)(g->high_max - g->high_min) + 1);
}

/* ADVERSARIAL: values on and next to the thresholds, extremes, negatives. */
KSA_HD int64_t ksa_gen_adversarial_value(const ksa_gen_params* g, uint64_t h) {
  uint32_t pick = (uint32_t)(h & 15u);
  int64_t j = (int64_t)((h >> 8) & 3u) - 1; /* -1..2 */
  switch (pick) {
    case 0: return g->low_max + j;
    case 1: return g->high_min + j;
    case 2: return INT64_MAX;
    case 3: return INT64_MIN;
    case 4: return -(int64_t)ksa_below(h, 1000);
    case 5: return INT64_MAX - (int64_t)ksa_below(h, 4);
    case 6: return 0;
    case 7: return (int64_t)(h >> 1);          /* any non-negative int64 */
    case 8: return (int64_t)h;                 /* any int64 */
    case 9: return (int64_t)0x100000000ll + j; /* just past 32 bits */
    case 10: return -(int64_t)0x100000000ll + j;
    default: return ksa_gen_bands_value(g, h);
  }
}

/* WALK: position after one step from `pos` (reflecting walls at 0 and high_max). */
KSA_HD int64_t ksa_gen_walk_step(const ksa_gen_params* g, int64_t pos, uint64_t h) {
  int64_t s = (int64_t)ksa_below(h, (uint64_t)(2 * g->walk_step + 1)) - g->walk_step;
  int64_t p = pos + s;
  if (p < 0) p = -p;
  if (p > g->high_max) p = 2 * g->high_max - p;
  if (p < 0) p = 0;
  return p;
}
KSA_HD int64_t ksa_gen_walk_start(const ksa_gen_params* g, uint64_t key) {
  return (int64_t)ksa_below(ksa_entry_word(key, 0), (uint64_t)g->high_max + 1);
}

/*
 * Split a total v >= 0 into A = 3 non-negative attribute values that sum to v (wrapping int64 sum
 * == v). With overflow_log2 != 0, one entry in 2^overflow_log2 is replaced by the triple
 * (INT64_MAX, 1 + v, 0), whose mathematically exact sum exceeds INT64_MAX and whose Go `int` sum
 * wraps to INT64_MIN + v (sqs/sqs.go:63).
 */
KSA_HD void ksa_gen_split3(const ksa_gen_params* g, int64_t v, uint64_t h, int64_t out[3]) {
  uint64_t h2 = ksa_mix64(h);
  if (g->overflow_log2 != 0 && (h2 >> (64 - g->overflow_log2)) == 0) {
    out[0] = INT64_MAX;
    out[1] = 1 + (v < 0 ? 0 : v);
    out[2] = 0;
    return;
  }
  if (v <= 0) { /* adversarial totals: keep the total in plane 0 */
    out[0] = v;
    out[1] = 0;
    out[2] = 0;
    return;
  }
  uint64_t uv = (uint64_t)v;
  uint64_t cap = uv > 0xFFFFFFFFull ? 0xFFFFFFFFull : uv; /* ksa_below wants range <= 2^32 */
  uint64_t p = ksa_below(h2, cap + 1);
  uint64_t rem = uv - p;
  uint64_t cap2 = rem > 0xFFFFFFFFull ? 0xFFFFFFFFull : rem;
  uint64_t q = ksa_below(h2 << 32, cap2 + 1);
  out[0] = (int64_t)p;
  out[1] = (int64_t)q;
  out[2] = (int64_t)(rem - q);
}

/* Value of a memoryless family (BANDS, ADVERSARIAL) for one entry word. */
KSA_HD int64_t ksa_gen_iid_value(const ksa_gen_params* g, uint64_t h) {
  return g->family == KSA_GEN_ADVERSARIAL ? ksa_gen_adversarial_value(g, h)
                                          : ksa_gen_bands_value(g, h);
}

/* Failure mask byte of one (trace, tick): bit k set with probability 2^-log2[k] (0 = never). The key
 * uses slot 0xF0, so the mask is independent of the depth of the same entry. */
KSA_HD uint8_t ksa_gen_fault_byte(uint64_t fault_key, uint64_t tick, uint32_t poll_log2,
                                  uint32_t up_log2, uint32_t down_log2) {
  const uint64_t h0 = ksa_entry_word(fault_key, tick);
  const uint64_t h1 = ksa_mix64(h0);
  const uint64_t h2 = ksa_mix64(h1);
  uint8_t f = 0;
  if (poll_log2 != 0 && (h0 >> (64 - poll_log2)) == 0) f |= KSA_FAIL_POLL;
  if (up_log2 != 0 && (h1 >> (64 - up_log2)) == 0) f |= KSA_FAIL_UP;
  if (down_log2 != 0 && (h2 >> (64 - down_log2)) == 0) f |= KSA_FAIL_DOWN;
  return f;
}
KSA_HD uint64_t ksa_fault_key(uint64_t seed, uint64_t trace) { return ksa_trace_key(seed, 0xF0u, trace); }

/* Initial replica count of a fresh trace: lo + hash(seed, trace) mod (hi - lo + 1). */
KSA_HD int32_t ksa_gen_replicas0(uint64_t seed, uint64_t trace, int32_t lo, int32_t hi) {
  uint64_t h = ksa_trace_key(seed, 0xA5u, trace);
  return (int32_t)(lo + (int64_t)ksa_below(h, (uint64_t)((int64_t)hi - lo) + 1));
}

#endif /* KSA_TRACEGEN_H */
