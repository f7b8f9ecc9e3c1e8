// ksa_ingest.cu — the step immediately before the path when the depths are recorded rather than
// synthetic: (1) trace-major [N][T] series -> the tick-major [T][A][ld] int64 array the replay
// kernels stream, (2) SQS attribute strings -> int64 with strconv.Atoi semantics, the decode the
// reference performs per attribute (sqs/sqs.go:58). Both are HBM-bound byte movers.
#include "ksa_internal.h"

#include <cstdlib>

namespace {

// Tiled transpose + widen. A block moves a tile of TT ticks x TN traces through shared memory.
// Load phase: each warp reads trace rows of the tile, 32 consecutive ticks per instruction (256 B for
// int64 sources), TT * 8 contiguous bytes per row. Store phase: each warp writes tick rows, 32
// consecutive traces per instruction, TN * 8 contiguous bytes per row. Every global access is a run
// of full 32-byte sectors; each element is read once and written once. The pitch TT + 1 (in 8-byte
// words, odd) makes the column reads of the store phase hit distinct banks.
// The tile shape trades DRAM row locality between the two sides: the kernel is latency/DRAM-page
// bound, not instruction bound (profiles/r01_ncu_transpose.txt), so longer contiguous runs help.
template <typename SRC, int TT, int TN, int THREADS>
__global__ void __launch_bounds__(THREADS) transpose_kernel(const SRC* __restrict__ src, long long src_ld,
                                                            long long T, long long N, int A, int a,
                                                            long long ld, long long* __restrict__ dst) {
  // [TN][TT + 1] in the source's element type: an int32 source keeps the tile at half the size,
  // which is what lets it use tiles twice as long in T (same 2 KB contiguous reads per trace row).
  extern __shared__ __align__(16) unsigned char tile_bytes[];
  SRC* const tile = reinterpret_cast<SRC*>(tile_bytes);
  constexpr int PITCH = TT + 1;
  constexpr int WARPS = THREADS / 32;
  const long long t0 = (long long)blockIdx.x * TT;
  const long long n0 = (long long)blockIdx.y * TN;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < TN / WARPS; ++k) {
    const int nl = warp + WARPS * k;
    const long long n = n0 + nl;
#pragma unroll
    for (int h = 0; h < TT / 32; ++h) {
      const int tl = lane + 32 * h;
      const long long t = t0 + tl;
      SRC v = 0;
      if (n < N && t < T) v = __ldg(src + n * src_ld + t);
      tile[nl * PITCH + tl] = v;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < TT / WARPS; ++k) {
    const int tl = warp + WARPS * k;
    const long long t = t0 + tl;
    if (t >= T) continue;
    long long* row = dst + (t * A + a) * ld;
#pragma unroll
    for (int h = 0; h < TN / 32; ++h) {
      const int nl = lane + 32 * h;
      const long long n = n0 + nl;
      if (n < N) __stcs(row + n, (long long)tile[nl * PITCH + tl]);  // sign-extends an int32 source
    }
  }
}

template <typename SRC, int TT, int TN, int THREADS>
int launch_transpose(const void* src, long long src_ld, long long T, long long N, int A, int a, long long ld,
                     long long* dst, cudaStream_t st) {
  const long long gx = (T + TT - 1) / TT;
  if (gx > 0x7fffffffll) return ksa_fail(KSA_ELIMIT, "T=%lld too large for one launch", T);
  constexpr size_t smem = sizeof(SRC) * (size_t)TN * (TT + 1);
  if (smem > 227 * 1024) return ksa_fail(KSA_EINVAL, "transpose shape %dx%d needs %zu bytes of shared memory", TT, TN, smem);
  auto kernel = transpose_kernel<SRC, TT, TN, THREADS>;
  if (smem > 48 * 1024) KSA_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // The trace axis is on gridDim.y (at most 65535 tiles): more traces go in slabs, one launch each,
  // so that every trace count the replay entry points accept can be ingested.
  constexpr long long kSlab = 65535ll * TN;
  for (long long nb = 0; nb < N; nb += kSlab) {
    const long long ns = N - nb < kSlab ? N - nb : kSlab;
    kernel<<<dim3((unsigned)gx, (unsigned)((ns + TN - 1) / TN)), THREADS, smem, st>>>(
        (const SRC*)src + nb * src_ld, src_ld, T, ns, A, a, ld, dst + nb);
    KSA_LAUNCH_CHECK("transpose_kernel");
  }
  return KSA_OK;
}

template <typename SRC>
int dispatch_transpose(int shape, const void* src, long long src_ld, long long T, long long N, int A, int a,
                       long long ld, long long* dst, cudaStream_t st) {
  switch (shape) {  // the two defaults in every build, the other tile shapes that were measured in the tuning build
    case 4: return launch_transpose<SRC, 64, 128, 512>(src, src_ld, T, N, A, a, ld, dst, st);
    case 7: return launch_transpose<SRC, 256, 32, 512>(src, src_ld, T, N, A, a, ld, dst, st);
#ifdef KSA_TUNING_VARIANTS
    case 0: return launch_transpose<SRC, 64, 64, 256>(src, src_ld, T, N, A, a, ld, dst, st);
    case 1: return launch_transpose<SRC, 128, 64, 512>(src, src_ld, T, N, A, a, ld, dst, st);
    case 2: return launch_transpose<SRC, 256, 32, 256>(src, src_ld, T, N, A, a, ld, dst, st);
    case 3: return launch_transpose<SRC, 128, 128, 1024>(src, src_ld, T, N, A, a, ld, dst, st);
    case 5: return launch_transpose<SRC, 128, 32, 256>(src, src_ld, T, N, A, a, ld, dst, st);
    case 6: return launch_transpose<SRC, 512, 32, 512>(src, src_ld, T, N, A, a, ld, dst, st);
    case 8: return launch_transpose<SRC, 512, 32, 1024>(src, src_ld, T, N, A, a, ld, dst, st);
    case 9: return launch_transpose<SRC, 256, 64, 512>(src, src_ld, T, N, A, a, ld, dst, st);
    case 10: return launch_transpose<SRC, 32, 256, 1024>(src, src_ld, T, N, A, a, ld, dst, st);
    case 11: return launch_transpose<SRC, 64, 256, 1024>(src, src_ld, T, N, A, a, ld, dst, st);
#endif
    default: return ksa_fail(KSA_EINVAL, "transpose shape %d is not in this build (KSA_TRANSPOSE_SHAPE)", shape);
  }
}
// profiles/r01_ingest_tune.txt: int64 source 256 ticks x 32 traces, 512 threads (1.00 of the copy
// bandwidth); int32 source 64 ticks x 128 traces, 512 threads (0.98; its tile is half the bytes).
constexpr int kDefaultTransposeShape = 7;
constexpr int kDefaultTransposeShapeInt32 = 4;

// strconv.Atoi on one fixed-width record. Go 1.15 semantics for base 10, 64-bit int:
//   ""            -> ErrSyntax          "+" / "-"    -> ErrSyntax
//   [+-]?[0-9]+   -> value, ErrRange if it does not fit int64 (|-2^63| allowed for '-')
//   anything else (spaces, '_', letters, a second sign) -> ErrSyntax
// The record ends at the first NUL or at `width` bytes.
// W > 0: compile-time width, fully unrolled so that the record stays in registers (no local-memory
// array, no divergent early exit: `live` masks the positions after the terminator). W == 0: run-time
// width, plain loop.
template <int W>
__device__ __forceinline__ bool atoi_record(const unsigned char* s, int width, long long* out) {
  const int n = W > 0 ? W : width;
  const unsigned c0 = s[0];
  const bool has_sign = (c0 == '+' || c0 == '-');
  const bool neg = c0 == '-';
  unsigned long long acc = 0;
  int digits = 0;
  bool bad = false, live = true;
#pragma unroll
  for (int p = 0; p < (W > 0 ? W : 32); ++p) {
    if (W == 0 && p >= n) break;
    const unsigned c = s[p];
    const unsigned d = c - '0';
    const bool skip = (p == 0) && has_sign;
    live = live && (c != 0);                     // the record ends at the first NUL
    const bool use = live && !skip;
    bad = bad || (use && d > 9u);                // ErrSyntax
    // acc * 10 + d must stay <= 2^63 (the signed limit is checked below): ErrRange otherwise
    bad = bad || (use && d <= 9u && acc > (0x8000000000000000ull - d) / 10ull);
    acc = (use && d <= 9u) ? acc * 10ull + d : acc;
    digits += (use && d <= 9u) ? 1 : 0;
  }
  if (bad || digits == 0) return false;          // "" or a bare sign is ErrSyntax too
  if (!neg && acc > 0x7fffffffffffffffull) return false;
  *out = neg ? (long long)(0ull - acc) : (long long)acc;
  return true;
}

// Fast path for records of at most 8 characters (sign included), i.e. every realistic queue depth:
// the 8 bytes are handled as one 64-bit word (SWAR). `x` holds the string little-endian (first
// character in the low byte), NUL padded. Returns false when the slow path must decide (a '+'/'-'
// sign, a non-digit, or an empty record); true with the value otherwise.
__device__ __forceinline__ bool atoi_swar8(unsigned long long x, long long* out) {
  // length = index of the first NUL byte (8 if none)
  const unsigned long long nul = (x - 0x0101010101010101ull) & ~x & 0x8080808080808080ull;
  const int len = nul ? (__ffsll((long long)nul) - 1) >> 3 : 8;
  if (len == 0) return false;
  // right-align: the string occupies the top `len` bytes, the low bytes become '0'
  const int sh = 8 * (8 - len);
  const unsigned long long xs = len == 8 ? x : ((x << sh) | (0x3030303030303030ull >> (8 * len)));
  // '0'..'9' -> 0..9; every other byte value maps to >= 10 under the XOR
  const unsigned long long d = xs ^ 0x3030303030303030ull;
  if (((d + 0x7676767676767676ull) | d) & 0x8080808080808080ull) return false;  // a non-digit (or sign)
  // 8 digits -> number: pairs, then 4-digit groups, then the two halves (most significant digit is
  // in the low byte)
  unsigned long long t = (d * 10ull + (d >> 8)) & 0x00FF00FF00FF00FFull;
  t = (t * 100ull + (t >> 16)) & 0x0000FFFF0000FFFFull;
  t = (t * 10000ull + (t >> 32)) & 0x00000000FFFFFFFFull;
  *out = (long long)t;
  return true;
}

// WIDTH_VEC = width / 16 when the records can be fetched as uint4, else 0. A block decodes
// 256 * U records: thread t takes records t, t + 256, ... of the block's range, so every load and
// store instruction of a warp is contiguous, and the U loads of a thread are all issued before the
// first one is used. At U = 1 a full SM has only 32 KB of loads in flight: 0.80 of the copy
// bandwidth on 2^28 records; U = 2: 0.94, U = 4: 0.99, U = 8: 1.05 (profiles/r01_ingest_tune.txt).
template <int WIDTH_VEC, int U>
__global__ void __launch_bounds__(256) decimal_kernel(const char* __restrict__ text, int width,
                                                      long long count, long long* __restrict__ values,
                                                      unsigned char* __restrict__ fail) {
  const long long base = (long long)blockIdx.x * (256 * U) + threadIdx.x;
  if constexpr (WIDTH_VEC > 0) {
    uint4 w[U][WIDTH_VEC];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = base + u * 256;
      if (i < count) {
        const uint4* p = reinterpret_cast<const uint4*>(text + i * width);
#pragma unroll
        for (int q = 0; q < WIDTH_VEC; ++q) w[u][q] = __ldg(p + q);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = base + u * 256;
      if (i >= count) continue;
      unsigned rest = w[u][0].z | w[u][0].w;
#pragma unroll
      for (int q = 1; q < WIDTH_VEC; ++q) rest |= w[u][q].x | w[u][q].y | w[u][q].z | w[u][q].w;
      const unsigned long long x0 = (unsigned long long)w[u][0].x | ((unsigned long long)w[u][0].y << 32);
      long long v = 0;
      bool ok;
      // bytes 8.. all NUL and the first 8 are plain digits: done; anything else takes the exact
      // character-by-character path (rare: signs, long numbers, malformed strings)
      if (rest == 0u && atoi_swar8(x0, &v)) {
        ok = true;
      } else {
        unsigned char rec[32];
#pragma unroll
        for (int q = 0; q < WIDTH_VEC; ++q) reinterpret_cast<uint4*>(rec)[q] = w[u][q];
        ok = atoi_record<WIDTH_VEC * 16>(rec, width, &v);
      }
      values[i] = ok ? v : 0;
      if (fail) fail[i] = ok ? 0 : KSA_FAIL_POLL;  // sqs.go:59-61 -> main.go:44-47
    }
  } else {
#pragma unroll 1
    for (int u = 0; u < U; ++u) {
      const long long i = base + u * 256;
      if (i >= count) continue;
      unsigned char rec[32];
      for (int q = 0; q < width; ++q) rec[q] = (unsigned char)__ldg(text + i * width + q);
      long long v = 0;
      const bool ok = atoi_record<0>(rec, width, &v);
      values[i] = ok ? v : 0;
      if (fail) fail[i] = ok ? 0 : KSA_FAIL_POLL;
    }
  }
}

// Failure bytes arrive trace-major next to the decoded values; the replay wants them tick-major.
// 64 x 64 byte tiles through shared memory; `accumulate` ORs into the destination, so the masks of
// several attributes (any failed Atoi fails the whole poll, sqs/sqs.go:58-61) add up.
__global__ void __launch_bounds__(256) transpose_mask_kernel(const unsigned char* __restrict__ src, long long src_ld,
                                                             long long T, long long N, long long ld,
                                                             unsigned char* __restrict__ dst, int accumulate) {
  __shared__ unsigned char tile[64][65];
  const long long t0 = (long long)blockIdx.x * 64, n0 = (long long)blockIdx.y * 64;
  const int lane = threadIdx.x & 63, row = threadIdx.x >> 6;  // 4 rows of 64 per pass
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int nl = row + 4 * k;
    const long long n = n0 + nl, t = t0 + lane;
    tile[nl][lane] = (n < N && t < T) ? src[n * src_ld + t] : (unsigned char)0;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int tl = row + 4 * k;
    const long long t = t0 + tl, n = n0 + lane;
    if (t < T && n < N) {
      unsigned char v = tile[lane][tl];
      if (accumulate) v |= dst[t * ld + n];
      dst[t * ld + n] = v;
    }
  }
}

}  // namespace

extern "C" int ksa_ingest_transpose(const void* src_dev, int32_t src_bytes, int64_t src_ld, int64_t T,
                                    int64_t N, int32_t A, int32_t a, int64_t ld, int64_t* depth_dev,
                                    void* stream) {
  if (!src_dev || !depth_dev) return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (src_bytes != 8 && src_bytes != 4) return ksa_fail(KSA_EINVAL, "src_bytes=%d: 8 (int64) or 4 (int32)", src_bytes);
  if (A < 1 || A > KSA_MAX_ATTRS || a < 0 || a >= A) return ksa_fail(KSA_EINVAL, "bad plane a=%d of A=%d", a, A);
  if (T < 0 || N < 0 || src_ld < T || ld < N) return ksa_fail(KSA_EINVAL, "bad T/N/src_ld/ld");
  if (T == 0 || N == 0) return KSA_OK;
  int shape = src_bytes == 8 ? kDefaultTransposeShape : kDefaultTransposeShapeInt32;
  if (const char* e = std::getenv("KSA_TRANSPOSE_SHAPE")) shape = std::atoi(e);  // tuning knob (tools/)
  cudaStream_t st = (cudaStream_t)stream;
  return src_bytes == 8
             ? dispatch_transpose<long long>(shape, src_dev, src_ld, T, N, A, a, ld, (long long*)depth_dev, st)
             : dispatch_transpose<int>(shape, src_dev, src_ld, T, N, A, a, ld, (long long*)depth_dev, st);
}

extern "C" int ksa_ingest_decimal(const char* text_dev, int32_t width, int64_t count, int64_t* values_dev,
                                  uint8_t* fail_dev, void* stream) {
  if (!text_dev || !values_dev) return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (width < 1 || width > 32) return ksa_fail(KSA_EINVAL, "width=%d outside [1,32]", width);
  if (count < 0) return ksa_fail(KSA_EINVAL, "count < 0");
  if (count == 0) return KSA_OK;
  int U = 8;
#ifdef KSA_TUNING_VARIANTS
  if (const char* e = std::getenv("KSA_DECIMAL_UNROLL")) U = std::atoi(e);  // tuning knob (tools/): 1, 2, 4 or 8
  if (U != 1 && U != 2 && U != 4) U = 8;
#endif
  const long long blocks = (count + 256 * U - 1) / (256 * U);
  if (blocks > 0x7fffffffll) return ksa_fail(KSA_ELIMIT, "count too large for one launch");
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (width % 16 == 0) && ksa_aligned(text_dev, 16);
  long long* out = (long long*)values_dev;
#define KSA_DEC(WV, UU) decimal_kernel<WV, UU><<<(unsigned)blocks, 256, 0, st>>>(text_dev, width, count, out, fail_dev)
#ifdef KSA_TUNING_VARIANTS
#define KSA_DEC_U(WV) \
  if (U == 1) KSA_DEC(WV, 1); else if (U == 2) KSA_DEC(WV, 2); else if (U == 4) KSA_DEC(WV, 4); else KSA_DEC(WV, 8)
#else
#define KSA_DEC_U(WV) KSA_DEC(WV, 8)  // profiles/r01_ingest_tune.txt: 8 records per thread
#endif
  if (vec && width == 16) {
    KSA_DEC_U(1);
  } else if (vec && width == 32) {
    KSA_DEC_U(2);
  } else {
    KSA_DEC_U(0);
  }
#undef KSA_DEC_U
#undef KSA_DEC
  KSA_LAUNCH_CHECK("decimal_kernel");
  return KSA_OK;
}

extern "C" int ksa_ingest_transpose_mask(const uint8_t* src_dev, int64_t src_ld, int64_t T, int64_t N, int64_t ld,
                                         uint8_t* fail_dev, int32_t accumulate, void* stream) {
  if (!src_dev || !fail_dev) return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (T < 0 || N < 0 || src_ld < T || ld < N) return ksa_fail(KSA_EINVAL, "bad T/N/src_ld/ld");
  if (T == 0 || N == 0) return KSA_OK;
  const long long gx = (T + 63) / 64;
  if (gx > 0x7fffffffll) return ksa_fail(KSA_ELIMIT, "T=%lld too large for one launch", (long long)T);
  constexpr long long kSlab = 65535ll * 64;  // gridDim.y holds 65535 tiles of 64 traces: slabs beyond that
  for (long long nb = 0; nb < N; nb += kSlab) {
    const long long ns = N - nb < kSlab ? N - nb : kSlab;
    transpose_mask_kernel<<<dim3((unsigned)gx, (unsigned)((ns + 63) / 64)), 256, 0, (cudaStream_t)stream>>>(
        src_dev + nb * src_ld, src_ld, T, ns, ld, fail_dev + nb, accumulate);
    KSA_LAUNCH_CHECK("transpose_mask_kernel");
  }
  return KSA_OK;
}
