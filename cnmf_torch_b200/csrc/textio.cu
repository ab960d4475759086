// Host-side text writer for the tab-separated tables of a consensus call (no device code in this file).
//
// The reference writes its usage and spectra tables with DataFrame.to_csv(sep="\t") (save_df_to_text,
// /root/reference/src/cnmf/cnmf.py:44-45); for a float frame pandas formats the values with numpy's shortest
// round-trip repr (values.astype(str)) and then walks the rows in Python.  At cfg-3 the usage table is 100 000 cells
// x k and that walk, once per K, was most of a consensus call.  Here: the same digits (std::to_chars = shortest
// round-trip, re-laid-out by numpy's rules: positional for 1e-4 <= |x| < 1e16 (float32: 1e6), else scientific with a two-digit
// exponent, "nan" -> empty field like pandas' na_rep), rows formatted by a few host threads, one fwrite per thread
// block.  tests/test_host_contract.py compares the bytes with pandas' on random bit patterns and the edge cases.
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

namespace cnmf {

// numpy's repr of a finite or non-finite float/double, appended to `out`
template <typename T>
static void append_repr(std::string& out, T v) {
  if (std::isnan(v)) return;                        // pandas: na_rep = ""
  if (std::isinf(v)) { out += v < 0 ? "-inf" : "inf"; return; }
  char buf[64];
  if (v == 0) { out += std::signbit(v) ? "-0.0" : "0.0"; return; }
  // shortest digits in scientific form: [-]d[.ddd]e[+-]XX
  auto r = std::to_chars(buf, buf + sizeof(buf) - 1, v, std::chars_format::scientific);
  *r.ptr = 0;
  const char* p = buf;
  const char* end = r.ptr;
  if (*p == '-') { out += '-'; ++p; }
  const char* e = p;
  while (e < end && *e != 'e') ++e;
  char digits[32];
  int nd = 0;
  for (const char* q = p; q < e; ++q)
    if (*q != '.') digits[nd++] = *q;
  const int exp10 = atoi(e + 1);
  const double a = std::fabs((double)v);
  const double upper = sizeof(T) == 4 ? 1e6 : 1e16;   // numpy's cut-over to scientific notation, per type
  if (a >= 1e-4 && a < upper) {
    if (exp10 >= 0) {                               // ddd[000].ddd | ddd.0
      const int int_digits = exp10 + 1;
      for (int i = 0; i < int_digits; ++i) out += i < nd ? digits[i] : '0';
      out += '.';
      if (nd > int_digits) out.append(digits + int_digits, nd - int_digits); else out += '0';
    } else {                                        // 0.000ddd
      out += "0.";
      for (int i = 0; i < -exp10 - 1; ++i) out += '0';
      out.append(digits, nd);
    }
  } else {                                          // d[.ddd]e[+-]XX (at least two exponent digits)
    out += digits[0];
    if (nd > 1) { out += '.'; out.append(digits + 1, nd - 1); }
    out.append(e, end - e);
  }
}

template <typename T>
static void format_rows(const T* data, int64_t ld, int64_t k, int64_t r0, int64_t r1, const char* index_blob,
                        const int64_t* index_off, std::string& out) {
  out.reserve((size_t)(r1 - r0) * (size_t)(16 + 12 * k));
  for (int64_t i = r0; i < r1; ++i) {
    out.append(index_blob + index_off[i], (size_t)(index_off[i + 1] - index_off[i]));
    const T* row = data + i * ld;
    for (int64_t j = 0; j < k; ++j) {
      out += '\t';
      append_repr(out, row[j]);
    }
    out += '\n';
  }
}

template <typename T>
static int write_tsv(const char* path, const char* header, int64_t header_len, const char* index_blob,
                     const int64_t* index_off, const T* data, int64_t n, int64_t k, int64_t ld, int threads) {
  FILE* f = fopen(path, "wb");
  if (!f) { set_error("write_tsv: cannot open %s", path); return CNMF_B200_ERR_INVALID; }
  bool ok = fwrite(header, 1, (size_t)header_len, f) == (size_t)header_len;
  const int64_t block = 8192;                       // rows per task; tasks are taken in order, written in order
  const int64_t ntask = (n + block - 1) / block;
  int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  if (nt > 16) nt = 16;
  if (nt > ntask) nt = (int)(ntask > 0 ? ntask : 1);
  for (int64_t t0 = 0; t0 < ntask && ok; t0 += nt) {
    const int m = (int)((ntask - t0) < nt ? (ntask - t0) : nt);
    std::vector<std::string> bufs(m);
    std::vector<std::thread> th;
    for (int w = 0; w < m; ++w) {
      const int64_t r0 = (t0 + w) * block, r1 = r0 + block < n ? r0 + block : n;
      th.emplace_back([&, w, r0, r1] { format_rows(data, ld, k, r0, r1, index_blob, index_off, bufs[w]); });
    }
    for (auto& x : th) x.join();
    for (int w = 0; w < m && ok; ++w) ok = fwrite(bufs[w].data(), 1, bufs[w].size(), f) == bufs[w].size();
  }
  ok = (fclose(f) == 0) && ok;
  if (!ok) { set_error("write_tsv: write to %s failed", path); return CNMF_B200_ERR_INVALID; }
  return 0;
}

}  // namespace cnmf

extern "C" int cnmf_b200_write_tsv(const char* path, const char* header, int64_t header_len, const char* index_blob,
                                   const int64_t* index_off, const void* data, int32_t is_f64, int64_t n, int64_t k,
                                   int64_t ld, int32_t threads) {
  CNMF_NVTX("cnmf_b200_write_tsv");
  using namespace cnmf;
  CNMF_CHECK_ARG(path && header && index_blob && index_off && data, "write_tsv: null pointer");
  CNMF_CHECK_ARG(n >= 0 && k >= 0 && ld >= k && header_len >= 0, "write_tsv: bad shape");
  if (is_f64) return write_tsv(path, header, header_len, index_blob, index_off, static_cast<const double*>(data), n, k, ld, threads);
  return write_tsv(path, header, header_len, index_blob, index_off, static_cast<const float*>(data), n, k, ld, threads);
}
