// Host side of the entropy stage: the reference's range coder (CompressAI 1.2.4
// cpp_exts/rans/rans_interface.cpp over ryg_rans rans64.h) and table construction
// (cpp_exts/ops/ops.cpp::pmf_to_quantized_cdf, EntropyBottleneck.update, GaussianConditional.update),
// restated in C++ because the package is not installable offline.
// The coder is sequential by construction (one 64-bit rANS state per string) and stays on the
// CPU; bench.py times it separately from the GPU hot path (north_star).
//
// Bitstream format (must not change): 64-bit state, L = 2^31, 32-bit words, 16-bit precision,
// 4-bit bypass escape; symbols are pushed in order and encoded in reverse; output = the written
// tail of the word buffer in native little-endian order.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <vector>

#include "../../include/pcc_b200.h"

namespace pcc {
void set_error(const char* fmt, ...);
}

namespace {

constexpr uint64_t kRansL = 1ull << 31;
constexpr uint32_t kPrecision = 16;
constexpr uint32_t kBypassBits = 4;
constexpr int32_t kMaxBypass = (1 << kBypassBits) - 1;

struct Table {
  const int32_t* cdf;
  int32_t max_value;
  int32_t offset;
};

// Per-(table, slot) encoder constants: ryg_rans Rans64EncSymbolInit.  x / freq is replaced by a
// 64x64->128 multiply with a rounded-up reciprocal; the result is bit-identical to the division
// form (that is the construction's guarantee), so the bitstream does not change.
struct EncSym {
  uint64_t x_max;
  uint64_t rcp_freq;
  uint32_t bias;
  uint32_t cmpl_freq;
  uint32_t rcp_shift;
  uint32_t freq;
};

inline EncSym make_enc_sym(uint32_t start, uint32_t freq) {
  EncSym s;
  s.freq = freq;
  s.x_max = ((kRansL >> kPrecision) << 32) * (uint64_t)freq;
  s.cmpl_freq = (1u << kPrecision) - freq;
  if (freq < 2) {
    s.rcp_freq = ~0ull;
    s.rcp_shift = 0;
    s.bias = start + (1u << kPrecision) - 1;
  } else {
    uint32_t shift = 0;
    while (freq > (1u << shift)) shift++;
    uint64_t x0 = freq - 1;
    const uint64_t x1 = 1ull << (shift + 31);
    const uint64_t t1 = x1 / freq;
    x0 += (x1 % freq) << 32;
    const uint64_t t0 = x0 / freq;
    s.rcp_freq = t0 + (t1 << 32);
    s.rcp_shift = shift - 1;
    s.bias = start;
  }
  return s;
}

// Encoder state writing 32-bit words back to front into a buffer that grows at the front.
struct BackWriter {
  std::vector<uint32_t> buf;
  size_t pos;  // index of the first written word
  uint64_t x;
  explicit BackWriter(size_t words) : buf(words), pos(words), x(kRansL) {}
  inline void emit(uint32_t w) {
    if (pos == 0) {
      const size_t old = buf.size();
      std::vector<uint32_t> nb(old * 2);
      memcpy(nb.data() + old, buf.data(), old * sizeof(uint32_t));
      buf.swap(nb);
      pos = old;
    }
    buf[--pos] = w;
  }
  inline void put(const EncSym& s) {
    if (x >= s.x_max) {
      emit((uint32_t)x);
      x >>= 32;
    }
    const uint64_t q = (uint64_t)(((unsigned __int128)x * s.rcp_freq) >> 64) >> s.rcp_shift;
    x = x + s.bias + q * s.cmpl_freq;
  }
  inline void put_bits(uint32_t val) {
    const uint64_t x_max = ((kRansL >> 16) << 32) * (uint64_t)(1u << (16 - kBypassBits));
    if (x >= x_max) {
      emit((uint32_t)x);
      x >>= 32;
    }
    x = (x << kBypassBits) | val;
  }
};

// maps a symbol to (table slot, escape payload); returns false when no escape is needed
inline bool classify(int32_t symbol, const Table& t, int32_t& slot, uint32_t& raw) {
  int32_t v = symbol - t.offset;
  raw = 0;
  if (v < 0) {
    raw = (uint32_t)(-2 * v - 1);
    v = t.max_value;
  } else if (v >= t.max_value) {
    raw = (uint32_t)(2 * (v - t.max_value));
    v = t.max_value;
  }
  slot = v;
  return v == t.max_value;
}

inline int nibbles_of(uint32_t raw) {
  int n = 0;
  while ((raw >> (n * kBypassBits)) != 0) ++n;
  return n;
}

double norm_ppf(double p) {
  // Acklam's rational approximation + two Halley refinements against erfc (|err| < 1e-15)
  static const double a[] = {-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02,
                             1.383577518672690e+02,  -3.066479806614716e+01, 2.506628277459239e+00};
  static const double b[] = {-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02,
                             6.680131188771972e+01,  -1.328068155288572e+01};
  static const double c[] = {-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00,
                             -2.549732539343734e+00, 4.374664141464968e+00,  2.938163982698783e+00};
  static const double d[] = {7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00,
                             3.754408661907416e+00};
  double x;
  if (p < 0.02425) {
    const double q = std::sqrt(-2 * std::log(p));
    x = (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
        ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1);
  } else if (p <= 1 - 0.02425) {
    const double q = p - 0.5, r = q * q;
    x = (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q /
        (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1);
  } else {
    const double q = std::sqrt(-2 * std::log(1 - p));
    x = -(((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
        ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1);
  }
  for (int it = 0; it < 2; ++it) {
    const double e = 0.5 * std::erfc(-x / std::sqrt(2.0)) - p;
    const double u = e * std::sqrt(2 * M_PI) * std::exp(x * x / 2);
    x = x - u / (1 + x * u / 2);
  }
  return x;
}

inline double std_cum(double v) { return 0.5 * std::erfc(-(1.0 / std::sqrt(2.0)) * v); }
inline double sigmoid(double v) { return 1.0 / (1.0 + std::exp(-v)); }
inline double softplus(double v) { return v > 30 ? v : std::log1p(std::exp(v)); }

// writes one padded cdf row from pmf[0..len) + tail
int write_cdf_row(const std::vector<float>& pmf, int len, float tail, int32_t* row, int stride) {
  std::vector<float> prob(pmf.begin(), pmf.begin() + len);
  prob.push_back(tail);
  std::vector<uint32_t> cdf(prob.size() + 1);
  const int rc = pcc_pmf_to_quantized_cdf(prob.data(), (int32_t)prob.size(), 16, cdf.data());
  if (rc) return rc;
  std::fill(row, row + stride, 0);
  for (size_t i = 0; i < cdf.size(); ++i) row[i] = (int32_t)cdf[i];
  return PCC_OK;
}

}  // namespace

extern "C" {

int64_t pcc_rans_encode_with_indexes(const int32_t* symbols, const int32_t* indexes, int64_t n,
                                     const int32_t* cdfs, int32_t cdf_stride, const int32_t* cdf_sizes,
                                     const int32_t* offsets, int32_t n_tables, uint8_t* out,
                                     int64_t out_cap) {
  if (n < 0 || (n > 0 && (!symbols || !indexes)) || !cdfs || !cdf_sizes || !offsets || !out || n_tables <= 0) {
    pcc::set_error("pcc_rans_encode_with_indexes: bad argument");
    return PCC_ERR_INVALID_ARG;
  }
  // encoder constants per table, built on first use
  std::vector<std::vector<EncSym>> enc((size_t)n_tables);
  BackWriter w((size_t)n / 2 + 1024);
  // Symbols are consumed in reverse.  Inside one symbol the reference pushed
  //   [table symbol][count digits ...][nibbles low-first]
  // so in reverse: nibbles high-first, count digits last-first, then the table symbol.
  for (int64_t i = n - 1; i >= 0; --i) {
    const int32_t idx = indexes[i];
    if (idx < 0 || idx >= n_tables) {
      pcc::set_error("pcc_rans_encode_with_indexes: index %d out of range at %lld", idx, (long long)i);
      return PCC_ERR_INVALID_ARG;
    }
    const Table t{cdfs + (int64_t)idx * cdf_stride, cdf_sizes[idx] - 2, offsets[idx]};
    std::vector<EncSym>& es = enc[idx];
    if (es.empty()) {
      es.resize((size_t)t.max_value + 1);
      for (int32_t v = 0; v <= t.max_value; ++v) {
        const int32_t f = t.cdf[v + 1] - t.cdf[v];
        if (f <= 0) {
          pcc::set_error("pcc_rans_encode_with_indexes: table %d has a non-positive frequency", idx);
          return PCC_ERR_INVALID_ARG;
        }
        es[v] = make_enc_sym((uint32_t)t.cdf[v], (uint32_t)f);
      }
    }
    int32_t slot;
    uint32_t raw;
    if (classify(symbols[i], t, slot, raw)) {
      const int nb = nibbles_of(raw);
      for (int j = nb - 1; j >= 0; --j) w.put_bits((raw >> (j * kBypassBits)) & kMaxBypass);
      const int full = nb / kMaxBypass;  // saturated count digits, pushed first by the reference
      w.put_bits((uint32_t)(nb - full * kMaxBypass));
      for (int j = 0; j < full; ++j) w.put_bits((uint32_t)kMaxBypass);
    }
    w.put(es[slot]);
  }
  w.emit((uint32_t)(w.x >> 32));
  w.emit((uint32_t)w.x);
  const int64_t nbytes = (int64_t)(w.buf.size() - w.pos) * 4;
  if (nbytes > out_cap) {
    pcc::set_error("pcc_rans_encode_with_indexes: output capacity %lld < %lld", (long long)out_cap,
                   (long long)nbytes);
    return PCC_ERR_CAPACITY;
  }
  memcpy(out, w.buf.data() + w.pos, (size_t)nbytes);
  return nbytes;
}

int pcc_rans_decode_with_indexes(const uint8_t* enc, int64_t nbytes, const int32_t* indexes, int64_t n,
                                 const int32_t* cdfs, int32_t cdf_stride, const int32_t* cdf_sizes,
                                 const int32_t* offsets, int32_t n_tables, int32_t* symbols) {
  if (!enc || nbytes < 8 || (n > 0 && (!indexes || !symbols)) || !cdfs || !cdf_sizes || !offsets) {
    pcc::set_error("pcc_rans_decode_with_indexes: bad argument");
    return PCC_ERR_INVALID_ARG;
  }
  const int64_t nwords = nbytes / 4;
  std::vector<uint32_t> words((size_t)nwords + 2, 0u);
  memcpy(words.data(), enc, (size_t)nwords * 4);
  const uint32_t* ptr = words.data();
  const uint32_t* end = words.data() + nwords;
  uint64_t x = (uint64_t)ptr[0] | ((uint64_t)ptr[1] << 32);
  ptr += 2;
  auto refill = [&]() {
    if (x < kRansL) {
      x = (x << 32) | (ptr < end ? *ptr : 0u);
      ++ptr;
    }
  };
  auto get_bits = [&]() -> int32_t {
    const int32_t v = (int32_t)(x & ((1u << kBypassBits) - 1));
    x >>= kBypassBits;
    refill();
    return v;
  };
  // per table: slot of the symbol containing cum = bucket << kLutShift, built on first use; the
  // reference finds the slot with a linear std::find_if over the row, which yields the same slot
  constexpr int kLutShift = 4;
  std::vector<std::vector<uint16_t>> lut((size_t)n_tables);
  for (int64_t i = 0; i < n; ++i) {
    const int32_t idx = indexes[i];
    if (idx < 0 || idx >= n_tables) {
      pcc::set_error("pcc_rans_decode_with_indexes: index %d out of range", idx);
      return PCC_ERR_INVALID_ARG;
    }
    const int32_t* cdf = cdfs + (int64_t)idx * cdf_stride;
    const int32_t size = cdf_sizes[idx];
    const int32_t max_value = size - 2;
    std::vector<uint16_t>& L = lut[idx];
    if (L.empty()) {
      L.resize((size_t)(1u << (kPrecision - kLutShift)));
      int32_t slot = 0;
      for (uint32_t b = 0; b < L.size(); ++b) {
        const int32_t c = (int32_t)(b << kLutShift);
        while (slot + 1 < size - 1 && cdf[slot + 1] <= c) ++slot;
        L[b] = (uint16_t)slot;
      }
    }
    const uint32_t cum = (uint32_t)(x & ((1u << kPrecision) - 1));
    int32_t s = L[cum >> kLutShift];
    while (cdf[s + 1] <= (int32_t)cum) ++s;  // first entry > cum, minus one
    const uint32_t start = (uint32_t)cdf[s], freq = (uint32_t)(cdf[s + 1] - cdf[s]);
    x = (uint64_t)freq * (x >> kPrecision) + cum - start;
    refill();
    int32_t value = s;
    if (value == max_value) {
      int32_t val = get_bits();
      int32_t nb = val;
      while (val == kMaxBypass) {
        val = get_bits();
        nb += val;
      }
      if (nb > 32 / kBypassBits) {  // a 32-bit value has at most 8 bypass nibbles: anything more is a corrupt stream
        pcc::set_error("pcc_rans_decode_with_indexes: corrupt stream (bypass length %d at symbol %lld)", nb, (long long)i);
        return PCC_ERR_INVALID_ARG;
      }
      int32_t raw = 0;
      for (int j = 0; j < nb; ++j) raw |= (int32_t)((uint32_t)get_bits() << (j * kBypassBits));
      value = raw >> 1;
      value = (raw & 1) ? (-value - 1) : (value + max_value);
    }
    symbols[i] = value + offsets[idx];
    if (ptr > end) {  // a valid stream is consumed exactly; reading past it means it was truncated or tampered with
      pcc::set_error("pcc_rans_decode_with_indexes: corrupt stream (exhausted at symbol %lld of %lld)", (long long)i,
                     (long long)n);
      return PCC_ERR_INVALID_ARG;
    }
  }
  return PCC_OK;
}

int pcc_pmf_to_quantized_cdf(const float* pmf, int32_t n, int32_t precision, uint32_t* cdf) {
  if (!pmf || !cdf || n <= 0 || precision < 1 || precision > 16) {
    pcc::set_error("pcc_pmf_to_quantized_cdf: bad argument");
    return PCC_ERR_INVALID_ARG;
  }
  for (int i = 0; i < n; ++i)
    if (!(pmf[i] >= 0) || !std::isfinite(pmf[i])) {
      pcc::set_error("pcc_pmf_to_quantized_cdf: invalid pmf value at %d", i);
      return PCC_ERR_INVALID_ARG;
    }
  const float scale = (float)(1 << precision);
  uint32_t total = 0;
  cdf[0] = 0;
  for (int i = 0; i < n; ++i) {
    cdf[i + 1] = (uint32_t)std::round(pmf[i] * scale);
    total += cdf[i + 1];
  }
  if (total == 0) {
    pcc::set_error("pcc_pmf_to_quantized_cdf: pmf sums to zero");
    return PCC_ERR_INVALID_ARG;
  }
  uint32_t run = 0;
  for (int i = 0; i <= n; ++i) {
    run += (uint32_t)((((uint64_t)1 << precision) * cdf[i]) / total);
    cdf[i] = run;
  }
  cdf[n] = 1u << precision;
  for (int i = 0; i < n; ++i) {
    if (cdf[i] != cdf[i + 1]) continue;
    uint32_t best = ~0u;
    int steal = -1;
    for (int j = 0; j < n; ++j) {
      const uint32_t f = cdf[j + 1] - cdf[j];
      if (f > 1 && f < best) {
        best = f;
        steal = j;
      }
    }
    if (steal < 0) {
      pcc::set_error("pcc_pmf_to_quantized_cdf: cannot make all frequencies positive");
      return PCC_ERR_INVALID_ARG;
    }
    if (steal < i)
      for (int j = steal + 1; j <= i; ++j) cdf[j]--;
    else
      for (int j = i + 1; j <= steal; ++j) cdf[j]++;
  }
  return PCC_OK;
}

int pcc_gaussian_cdf_tables(const float* scale_table, int32_t n, double tail_mass, int32_t* cdf,
                            int32_t* cdf_stride, int32_t* cdf_length, int32_t* offset) {
  if (!scale_table || n <= 0 || !cdf_stride) {
    pcc::set_error("pcc_gaussian_cdf_tables: bad argument");
    return PCC_ERR_INVALID_ARG;
  }
  const float mult = (float)(-norm_ppf(tail_mass / 2));
  std::vector<int32_t> center(n);
  int32_t max_len = 0;
  for (int i = 0; i < n; ++i) {
    center[i] = (int32_t)std::ceil(scale_table[i] * mult);  // float32 product, as torch
    max_len = std::max(max_len, 2 * center[i] + 1);
  }
  if (!cdf) {
    *cdf_stride = max_len + 2;
    return PCC_OK;
  }
  if (*cdf_stride < max_len + 2 || !cdf_length || !offset) {
    pcc::set_error("pcc_gaussian_cdf_tables: cdf_stride too small");
    return PCC_ERR_CAPACITY;
  }
  std::vector<float> pmf(max_len);
  for (int i = 0; i < n; ++i) {
    const double s = (double)scale_table[i];
    const int len = 2 * center[i] + 1;
    float tail = 0.f;
    for (int j = 0; j < max_len; ++j) {
      const double v = std::fabs((double)(j - center[i]));
      const double upper = std_cum((0.5 - v) / s), lower = std_cum((-0.5 - v) / s);
      pmf[j] = (float)(upper - lower);
      if (j == 0) tail = (float)(2 * lower);
    }
    const int rc = write_cdf_row(pmf, len, tail, cdf + (int64_t)i * *cdf_stride, *cdf_stride);
    if (rc) return rc;
    cdf_length[i] = len + 2;
    offset[i] = -center[i];
  }
  return PCC_OK;
}

int pcc_factorized_cdf_tables(const float* const* matrices, const float* const* biases,
                              const float* const* factors, const float* quantiles, int32_t c,
                              int32_t* cdf, int32_t* cdf_stride, int32_t* cdf_length, int32_t* offset) {
  if (!matrices || !biases || !factors || !quantiles || c <= 0 || !cdf_stride) {
    pcc::set_error("pcc_factorized_cdf_tables: bad argument");
    return PCC_ERR_INVALID_ARG;
  }
  static const int f[6] = {1, 3, 3, 3, 3, 1};
  std::vector<int32_t> minima(c), maxima(c);
  int32_t max_len = 0;
  for (int ch = 0; ch < c; ++ch) {
    const float q0 = quantiles[3 * ch], med = quantiles[3 * ch + 1], q2 = quantiles[3 * ch + 2];
    minima[ch] = std::max(0, (int32_t)std::ceil(med - q0));
    maxima[ch] = std::max(0, (int32_t)std::ceil(q2 - med));
    max_len = std::max(max_len, minima[ch] + maxima[ch] + 1);
  }
  if (!cdf) {
    *cdf_stride = max_len + 2;
    return PCC_OK;
  }
  if (*cdf_stride < max_len + 2 || !cdf_length || !offset) {
    pcc::set_error("pcc_factorized_cdf_tables: cdf_stride too small");
    return PCC_ERR_CAPACITY;
  }
  auto logits = [&](int ch, double v) {
    double h[3] = {v, 0, 0}, g[3];
    for (int l = 0; l < 5; ++l) {
      const int fo = f[l + 1], fi = f[l];
      const float* M = matrices[l] + (int64_t)ch * fo * fi;
      const float* B = biases[l] + (int64_t)ch * fo;
      for (int j = 0; j < fo; ++j) {
        double t = (double)B[j];
        for (int k = 0; k < fi; ++k) t += softplus((double)M[j * fi + k]) * h[k];
        if (l < 4) t += std::tanh((double)factors[l][(int64_t)ch * fo + j]) * std::tanh(t);
        g[j] = t;
      }
      for (int j = 0; j < fo; ++j) h[j] = g[j];
    }
    return h[0];
  };
  std::vector<float> pmf(max_len);
  for (int ch = 0; ch < c; ++ch) {
    const float med = quantiles[3 * ch + 1];
    const float pmf_start = med - (float)minima[ch];
    const int len = minima[ch] + maxima[ch] + 1;
    float tail = 0.f;
    double first_lower = 0, last_upper = 0;
    for (int j = 0; j < max_len; ++j) {
      const double s = (double)((float)j + pmf_start);  // float32 sample grid, as torch
      const double lower = logits(ch, s - 0.5), upper = logits(ch, s + 0.5);
      const double su = lower + upper;
      const double sign = su > 0 ? -1.0 : (su < 0 ? 1.0 : 0.0);
      pmf[j] = (float)std::fabs(sigmoid(sign * upper) - sigmoid(sign * lower));
      if (j == 0) first_lower = lower;
      if (j == max_len - 1) last_upper = upper;
    }
    tail = (float)(sigmoid(first_lower) + sigmoid(-last_upper));
    const int rc = write_cdf_row(pmf, len, tail, cdf + (int64_t)ch * *cdf_stride, *cdf_stride);
    if (rc) return rc;
    cdf_length[ch] = len + 2;
    offset[ch] = -minima[ch];
  }
  return PCC_OK;
}

}  // extern "C"
