// ORACLE — TEST INFRASTRUCTURE ONLY.  **parity unpinned** (see below).
//
// CPU restatement of the numeric primitives underneath the reference's encode/decode hot
// path (ColorModel.compress / ColorModel.decompress, /root/reference/model/model.py:95-208).
// The reference delegates every one of these primitives to MinkowskiEngine 0.5.x (unpinned,
// not in requirements.txt) and CompressAI 1.2.4 (requirements.txt:9); neither package is
// present in /root/reference or installable offline, so what follows restates their
// *published* algorithms and anchors on the reference's call
// sites, which are cited per function.
//
// "parity unpinned" at the third-party boundary: the reference holds no golden vectors,
// known-answer tests or fixtures for this path and MinkowskiEngine /
// CompressAI cannot be run here.  The pins that do exist are created by this repo:
//   * tests/test_oracle_pins.py  — sparse conv / strided / generative-transposed conv ==
//                                  torch.nn.functional.conv3d / conv_transpose3d identities,
//                                  hand-computed rANS known answers + big-int restatement,
//                                  Gaussian tables against scipy.stats.norm
//   * tests/golden/make_golden.py + tests/test_golden.py — the reference's OWN model/*.py and
//                                  utils.py, imported unmodified from /root/reference and executed
//                                  over these primitives (oracle/ref_shims.py); fixtures committed.
//                                  Pins control flow and state_dict names, not ME/CompressAI internals.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load this library.  The product (csrc/) never links or calls it.
//
// Build: see oracle/Makefile (g++ -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp).
//
// Floating-point contract: for every output row o and output channel co
//     acc = +0.0f
//     for k = 0 .. K-1 (kernel offsets, ascending), if neighbour row i = nbr[k][o] exists:
//         for ci = 0 .. Cin-1 (ascending):  acc = fmaf(in[i][ci], W[k][ci][co], acc)
//     out[o][co] = acc (+ bias[co], one rounded add, when bias is given)
// A missing neighbour is skipped; feeding a zero row instead is bit-identical because the
// accumulator starts at +0 and can never become -0 under round-to-nearest.

#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>
#include <immintrin.h>
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {

int orc_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

// ---------------------------------------------------------------------------------------
// Coordinate keys.  A coordinate row (b, x, y, z) — batch first, as ME.SparseTensor.C
// (model/model.py:117-122) — is packed so that unsigned key order == lexicographic
// (b, x, y, z) order, which is the canonical order of utils.sort_tensor / sort_points
// (/root/reference/utils.py:155-204).
// ---------------------------------------------------------------------------------------
static inline uint64_t pack_key(int32_t b, int32_t x, int32_t y, int32_t z) {
  return ((uint64_t)(uint16_t)b << 48) | ((uint64_t)(uint16_t)(x + 32768) << 32) |
         ((uint64_t)(uint16_t)(y + 32768) << 16) | (uint64_t)(uint16_t)(z + 32768);
}

void orc_pack_keys(const int32_t* coords, int64_t n, uint64_t* keys) {
  for (int64_t i = 0; i < n; ++i)
    keys[i] = pack_key(coords[4 * i], coords[4 * i + 1], coords[4 * i + 2], coords[4 * i + 3]);
}

void orc_unpack_keys(const uint64_t* keys, int64_t n, int32_t* coords) {
  for (int64_t i = 0; i < n; ++i) {
    uint64_t k = keys[i];
    coords[4 * i + 0] = (int32_t)(k >> 48);
    coords[4 * i + 1] = (int32_t)((k >> 32) & 0xFFFF) - 32768;
    coords[4 * i + 2] = (int32_t)((k >> 16) & 0xFFFF) - 32768;
    coords[4 * i + 3] = (int32_t)(k & 0xFFFF) - 32768;
  }
}

// ---------------------------------------------------------------------------------------
// Kernel map (ME kernel_map.cuh [upstream-recalled]).
// For every output coordinate c_o and kernel offset k: nbr[k][o] = row of the input
// coordinate c_o + offs[k] (offsets already scaled by the tensor stride and, for transposed
// convolution, negated by the caller), or -1.  in_keys must be sorted ascending (canonical
// order) — lookup is a binary search, deliberately a different algorithm from the product's
// GPU hash table.
// Call sites this serves: every ME.MinkowskiConvolution / ...Transpose in model/transforms.py,
// model/blocks.py, model/entropy_models.py.
// ---------------------------------------------------------------------------------------
void orc_kernel_map(const uint64_t* in_keys, int64_t n_in, const int32_t* out_coords,
                    int64_t n_out, const int32_t* offs, int K, int32_t* nbr) {
#pragma omp parallel for schedule(static)
  for (int64_t o = 0; o < n_out; ++o) {
    const int32_t b = out_coords[4 * o], x = out_coords[4 * o + 1], y = out_coords[4 * o + 2],
                  z = out_coords[4 * o + 3];
    for (int k = 0; k < K; ++k) {
      int32_t qx = x + offs[3 * k], qy = y + offs[3 * k + 1], qz = z + offs[3 * k + 2];
      int32_t r = -1;
      if (qx >= -32768 && qx < 32768 && qy >= -32768 && qy < 32768 && qz >= -32768 && qz < 32768) {
        uint64_t key = pack_key(b, qx, qy, qz);
        const uint64_t* it = std::lower_bound(in_keys, in_keys + n_in, key);
        if (it != in_keys + n_in && *it == key) r = (int32_t)(it - in_keys);
      }
      nbr[(int64_t)k * n_out + o] = r;
    }
  }
}

// Row lookup of query coordinates in a sorted key set (SparseTensor.features_at_coordinates
// on on-grid integer queries == gather-or-zero; call sites
// model/blocks.py:37,50, model/transforms.py:96,124,262, model/entropy_models.py:364,401).
void orc_lookup(const uint64_t* in_keys, int64_t n_in, const uint64_t* q_keys, int64_t n_q,
                int32_t* rows) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n_q; ++i) {
    const uint64_t* it = std::lower_bound(in_keys, in_keys + n_in, q_keys[i]);
    rows[i] = (it != in_keys + n_in && *it == q_keys[i]) ? (int32_t)(it - in_keys) : -1;
  }
}

// ---------------------------------------------------------------------------------------
// Sparse convolution forward in the exact fp32 order stated in the header.
// Serves ME.MinkowskiConvolution (57 sites), ME.MinkowskiGenerativeConvolutionTranspose
// (model/blocks.py:84, model/entropy_models.py:286,290) and ME.MinkowskiConvolutionTranspose
// (model/entropy_models.py:298,302): they differ only in how nbr was built.
//   in  [n_in, cin]   W [K, cin, cout] (ME `kernel` layout)   bias [cout] or NULL
//   nbr [K, n_out]    out [n_out, cout]
// ---------------------------------------------------------------------------------------
static void conv_rows_scalar(const float* in, int cin, const int32_t* nbr, int64_t n_out, int K,
                             const float* W, int cout, const float* bias, float* out, int64_t o,
                             int c_begin, int c_end) {
  for (int co = c_begin; co < c_end; ++co) {
    float acc = 0.0f;
    for (int k = 0; k < K; ++k) {
      int32_t i = nbr[(int64_t)k * n_out + o];
      if (i < 0) continue;
      const float* a = in + (int64_t)i * cin;
      const float* w = W + (int64_t)k * cin * cout + co;
      for (int ci = 0; ci < cin; ++ci) acc = __builtin_fmaf(a[ci], w[(int64_t)ci * cout], acc);
    }
    if (bias) acc = acc + bias[co];
    out[o * cout + co] = acc;
  }
}

void orc_conv(const float* in, int64_t n_in, int cin, const int32_t* nbr, int64_t n_out, int K,
              const float* W, int cout, const float* bias, float* out) {
  (void)n_in;
  constexpr int R = 6;  // output rows per register tile
  std::vector<float> zero_row((size_t)cin, 0.0f);
  const float* zr = zero_row.data();
  const int cvec = cout & ~15;  // channels handled by the 16-wide AVX2/FMA path
#pragma omp parallel for schedule(dynamic, 8)
  for (int64_t o0 = 0; o0 < n_out; o0 += R) {
    const int rows = (int)std::min<int64_t>(R, n_out - o0);
    for (int c0 = 0; c0 < cvec; c0 += 16) {
      __m256 acc[R][2];
      for (int r = 0; r < R; ++r) acc[r][0] = acc[r][1] = _mm256_setzero_ps();
      for (int k = 0; k < K; ++k) {
        const float* a[R];
        bool any = false;
        for (int r = 0; r < R; ++r) {
          int32_t i = (r < rows) ? nbr[(int64_t)k * n_out + o0 + r] : -1;
          a[r] = (i >= 0) ? in + (int64_t)i * cin : zr;
          any |= (i >= 0);
        }
        if (!any) continue;
        const float* w = W + (int64_t)k * cin * cout + c0;
        for (int ci = 0; ci < cin; ++ci) {
          const __m256 w0 = _mm256_loadu_ps(w + (int64_t)ci * cout);
          const __m256 w1 = _mm256_loadu_ps(w + (int64_t)ci * cout + 8);
          for (int r = 0; r < R; ++r) {
            const __m256 av = _mm256_broadcast_ss(a[r] + ci);
            acc[r][0] = _mm256_fmadd_ps(av, w0, acc[r][0]);
            acc[r][1] = _mm256_fmadd_ps(av, w1, acc[r][1]);
          }
        }
      }
      for (int r = 0; r < rows; ++r) {
        __m256 v0 = acc[r][0], v1 = acc[r][1];
        if (bias) {
          v0 = _mm256_add_ps(v0, _mm256_loadu_ps(bias + c0));
          v1 = _mm256_add_ps(v1, _mm256_loadu_ps(bias + c0 + 8));
        }
        _mm256_storeu_ps(out + (o0 + r) * cout + c0, v0);
        _mm256_storeu_ps(out + (o0 + r) * cout + c0 + 8, v1);
      }
    }
    if (cvec < cout)
      for (int r = 0; r < rows; ++r)
        conv_rows_scalar(in, cin, nbr, n_out, K, W, cout, bias, out, o0 + r, cvec, cout);
  }
}

// Plain scalar version of the same contract; used by tests to check the vector path.
void orc_conv_scalar(const float* in, int64_t n_in, int cin, const int32_t* nbr, int64_t n_out,
                     int K, const float* W, int cout, const float* bias, float* out) {
  (void)n_in;
#pragma omp parallel for schedule(static)
  for (int64_t o = 0; o < n_out; ++o)
    conv_rows_scalar(in, cin, nbr, n_out, K, W, cout, bias, out, o, 0, cout);
}

// ---------------------------------------------------------------------------------------
// CompressAI cpp_exts/ops/ops.cpp::pmf_to_quantized_cdf [upstream-recalled, v1.2.4]
// (SURVEY.md Appendix B4).  Reached from ColorModel.update (model/model.py:30-36).
// Returns 0 on success, <0 on invalid input.  cdf has n+1 entries.
// ---------------------------------------------------------------------------------------
int orc_pmf_to_quantized_cdf(const float* pmf, int n, int precision, uint32_t* cdf) {
  for (int i = 0; i < n; ++i)
    if (pmf[i] < 0 || !std::isfinite(pmf[i])) return -1;
  cdf[0] = 0;
  for (int i = 0; i < n; ++i) cdf[i + 1] = (uint32_t)std::round(pmf[i] * (float)(1 << precision));
  uint32_t total = 0;
  for (int i = 0; i <= n; ++i) total += cdf[i];
  if (total == 0) return -2;
  for (int i = 0; i <= n; ++i)
    cdf[i] = (uint32_t)((((uint64_t)1 << precision) * (uint64_t)cdf[i]) / total);
  for (int i = 1; i <= n; ++i) cdf[i] += cdf[i - 1];
  cdf[n] = 1u << precision;
  for (int i = 0; i < n; ++i) {
    if (cdf[i] == cdf[i + 1]) {
      uint32_t best_freq = ~0u;
      int best_steal = -1;
      for (int j = 0; j < n; ++j) {
        uint32_t freq = cdf[j + 1] - cdf[j];
        if (freq > 1 && freq < best_freq) {
          best_freq = freq;
          best_steal = j;
        }
      }
      if (best_steal == -1) return -3;
      if (best_steal < i) {
        for (int j = best_steal + 1; j <= i; ++j) cdf[j]--;
      } else {
        for (int j = i + 1; j <= best_steal; ++j) cdf[j]++;
      }
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------------------
// rANS coder: CompressAI cpp_exts/rans/rans_interface.cpp over ryg_rans rans64.h
// [upstream-recalled].  Reached from EntropyBottleneck.compress /
// .decompress and GaussianConditional.compress / .decompress at
// model/entropy_models.py:352-353,372,393,408.
//   cdfs: [n_tables, cdf_stride] int32 (row-padded), cdf_sizes/offsets: [n_tables]
// ---------------------------------------------------------------------------------------
static const uint64_t RANS64_L = 1ull << 31;
static const int PRECISION = 16;
static const int BYPASS_PRECISION = 4;
static const int MAX_BYPASS_VAL = (1 << BYPASS_PRECISION) - 1;

struct OrcSym {
  uint16_t start;
  uint16_t range;
  bool bypass;
};

static inline void enc_put(uint64_t& x, uint32_t*& ptr, uint32_t start, uint32_t freq,
                           uint32_t scale_bits) {
  uint64_t x_max = ((RANS64_L >> scale_bits) << 32) * freq;
  if (x >= x_max) {
    ptr -= 1;
    *ptr = (uint32_t)x;
    x >>= 32;
  }
  x = ((x / freq) << scale_bits) + (x % freq) + start;
}

static inline void enc_put_bits(uint64_t& x, uint32_t*& ptr, uint32_t val, uint32_t nbits) {
  uint32_t freq = 1u << (16 - nbits);
  uint64_t x_max = ((RANS64_L >> 16) << 32) * freq;
  if (x >= x_max) {
    ptr -= 1;
    *ptr = (uint32_t)x;
    x >>= 32;
  }
  x = (x << nbits) | val;
}

// Returns number of bytes written to out (the caller passes out_cap bytes), or <0 on error
// (-1: capacity, -2: bad index/symbol).  A generous capacity is 8*(n + extra) + 16.
int64_t orc_rans_encode(const int32_t* symbols, const int32_t* indexes, int64_t n,
                        const int32_t* cdfs, int cdf_stride, const int32_t* cdf_sizes,
                        const int32_t* offsets, int n_tables, uint8_t* out, int64_t out_cap) {
  std::vector<OrcSym> syms;
  syms.reserve((size_t)n + 16);
  for (int64_t i = 0; i < n; ++i) {
    const int32_t idx = indexes[i];
    if (idx < 0 || idx >= n_tables) return -2;
    const int32_t* cdf = cdfs + (int64_t)idx * cdf_stride;
    const int32_t max_value = cdf_sizes[idx] - 2;
    int32_t value = symbols[i] - offsets[idx];
    uint32_t raw_val = 0;
    if (value < 0) {
      raw_val = (uint32_t)(-2 * value - 1);
      value = max_value;
    } else if (value >= max_value) {
      raw_val = (uint32_t)(2 * (value - max_value));
      value = max_value;
    }
    syms.push_back({(uint16_t)cdf[value], (uint16_t)(cdf[value + 1] - cdf[value]), false});
    if (value == max_value) {
      int32_t n_bypass = 0;
      while ((raw_val >> (n_bypass * BYPASS_PRECISION)) != 0) ++n_bypass;
      int32_t val = n_bypass;
      while (val >= MAX_BYPASS_VAL) {
        syms.push_back({(uint16_t)MAX_BYPASS_VAL, (uint16_t)(MAX_BYPASS_VAL + 1), true});
        val -= MAX_BYPASS_VAL;
      }
      syms.push_back({(uint16_t)val, (uint16_t)(val + 1), true});
      for (int32_t j = 0; j < n_bypass; ++j) {
        const int32_t v = (raw_val >> (j * BYPASS_PRECISION)) & MAX_BYPASS_VAL;
        syms.push_back({(uint16_t)v, (uint16_t)(v + 1), true});
      }
    }
  }
  // flush(): symbols are consumed in reverse, words are written back-to-front
  std::vector<uint32_t> buf(syms.size() + 2, 0xCC);
  uint32_t* ptr = buf.data() + buf.size();
  uint64_t x = RANS64_L;
  for (size_t i = syms.size(); i-- > 0;) {
    const OrcSym& s = syms[i];
    if (!s.bypass)
      enc_put(x, ptr, s.start, s.range, PRECISION);
    else
      enc_put_bits(x, ptr, s.start, BYPASS_PRECISION);
  }
  ptr -= 2;
  ptr[0] = (uint32_t)(x >> 0);
  ptr[1] = (uint32_t)(x >> 32);
  const int64_t nbytes = (int64_t)((buf.data() + buf.size()) - ptr) * 4;
  if (nbytes > out_cap) return -1;
  std::memcpy(out, ptr, (size_t)nbytes);
  return nbytes;
}

static inline uint32_t dec_get_bits(uint64_t& x, const uint32_t*& ptr, uint32_t n_bits) {
  uint32_t val = (uint32_t)(x & ((1u << n_bits) - 1));
  x >>= n_bits;
  if (x < RANS64_L) {
    x = (x << 32) | *ptr;
    ptr += 1;
  }
  return val;
}

int orc_rans_decode(const uint8_t* enc, int64_t nbytes, const int32_t* indexes, int64_t n,
                    const int32_t* cdfs, int cdf_stride, const int32_t* cdf_sizes,
                    const int32_t* offsets, int n_tables, int32_t* out) {
  if (nbytes < 8) return -1;
  std::vector<uint32_t> words((size_t)(nbytes / 4) + 4, 0);
  std::memcpy(words.data(), enc, (size_t)nbytes);
  const uint32_t* ptr = words.data();
  uint64_t x = (uint64_t)ptr[0] | ((uint64_t)ptr[1] << 32);
  ptr += 2;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t idx = indexes[i];
    if (idx < 0 || idx >= n_tables) return -2;
    const int32_t* cdf = cdfs + (int64_t)idx * cdf_stride;
    const int32_t size = cdf_sizes[idx];
    const int32_t max_value = size - 2;
    const uint32_t cum_freq = (uint32_t)(x & ((1u << PRECISION) - 1));
    int j = 0;  // std::find_if: first entry strictly greater than cum_freq
    while (j < size && !((uint32_t)cdf[j] > cum_freq)) ++j;
    const uint32_t s = (uint32_t)(j - 1);
    const uint32_t start = (uint32_t)cdf[s], freq = (uint32_t)(cdf[s + 1] - cdf[s]);
    x = (uint64_t)freq * (x >> PRECISION) + (x & ((1ull << PRECISION) - 1)) - start;
    if (x < RANS64_L) {
      x = (x << 32) | *ptr;
      ptr += 1;
    }
    int32_t value = (int32_t)s;
    if (value == max_value) {
      int32_t val = (int32_t)dec_get_bits(x, ptr, BYPASS_PRECISION);
      int32_t n_bypass = val;
      while (val == MAX_BYPASS_VAL) {
        val = (int32_t)dec_get_bits(x, ptr, BYPASS_PRECISION);
        n_bypass += val;
      }
      int32_t raw_val = 0;
      for (int k = 0; k < n_bypass; ++k) {
        val = (int32_t)dec_get_bits(x, ptr, BYPASS_PRECISION);
        raw_val |= val << (k * BYPASS_PRECISION);
      }
      value = raw_val >> 1;
      if (raw_val & 1)
        value = -value - 1;
      else
        value += max_value;
    }
    out[i] = value + offsets[idx];
  }
  return 0;
}

}  // extern "C"
