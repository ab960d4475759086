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

#include <zlib.h>

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

// ---- .df.npz replicate files --------------------------------------------------------------------------------------
// One archive = three deflated members (data.npy, index.npy, columns.npy), the layout np.savez_compressed gives the
// reference's save_df_to_npz (cnmf.py:32-34) and np.load reads back.  The Python writer this replaces spent ~0.5 ms
// of interpreter time per file under the GIL: 0.3 s for the 600 replicate files of cfg-2, most of factorize's wall time.
struct ZipMember {
  const char* name;
  uint32_t crc;
  uint32_t raw_len;
  std::string comp;          // raw deflate stream (or empty when `shared` is used)
  const char* shared;        // pre-deflated bytes owned by the caller
  uint32_t shared_len;
};

static bool deflate_raw(const unsigned char* a, size_t na, const unsigned char* b, size_t nb, int level, std::string& out,
                        uint32_t& crc) {
  z_stream zs;
  memset(&zs, 0, sizeof(zs));
  if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) return false;
  out.resize(deflateBound(&zs, (uLong)(na + nb)) + 64);
  zs.next_out = reinterpret_cast<Bytef*>(&out[0]);
  zs.avail_out = (uInt)out.size();
  crc = (uint32_t)crc32(0L, Z_NULL, 0);
  zs.next_in = const_cast<Bytef*>(a);
  zs.avail_in = (uInt)na;
  crc = (uint32_t)crc32(crc, a, (uInt)na);
  const int r0 = deflate(&zs, nb > 0 ? Z_NO_FLUSH : Z_FINISH);
  bool ok = nb > 0 ? r0 == Z_OK : r0 == Z_STREAM_END;
  if (nb > 0) {
    zs.next_in = const_cast<Bytef*>(b);
    zs.avail_in = (uInt)nb;
    crc = (uint32_t)crc32(crc, b, (uInt)nb);
    ok = ok && deflate(&zs, Z_FINISH) == Z_STREAM_END;
  }
  out.resize(zs.total_out);
  deflateEnd(&zs);
  return ok;
}

static void put16(std::string& s, uint32_t v) { s.push_back((char)(v & 255)); s.push_back((char)((v >> 8) & 255)); }
static void put32(std::string& s, uint32_t v) { put16(s, v & 0xFFFF); put16(s, v >> 16); }

static bool write_zip_file(const char* path, const ZipMember* m, int nm) {
  std::string out, central;
  for (int i = 0; i < nm; ++i) {
    const uint32_t clen = m[i].shared ? m[i].shared_len : (uint32_t)m[i].comp.size();
    const uint32_t nlen = (uint32_t)strlen(m[i].name);
    const uint32_t offset = (uint32_t)out.size();
    put32(out, 0x04034B50); put16(out, 20); put16(out, 0); put16(out, 8); put16(out, 0); put16(out, 0x21);
    put32(out, m[i].crc); put32(out, clen); put32(out, m[i].raw_len); put16(out, nlen); put16(out, 0);
    out.append(m[i].name, nlen);
    if (m[i].shared) out.append(m[i].shared, clen); else out.append(m[i].comp);
    put32(central, 0x02014B50); put16(central, 20); put16(central, 20); put16(central, 0); put16(central, 8);
    put16(central, 0); put16(central, 0x21); put32(central, m[i].crc); put32(central, clen); put32(central, m[i].raw_len);
    put16(central, nlen); put16(central, 0); put16(central, 0); put16(central, 0); put16(central, 0); put32(central, 0);
    put32(central, offset);
    central.append(m[i].name, nlen);
  }
  const uint32_t cd_off = (uint32_t)out.size();
  out.append(central);
  put32(out, 0x06054B50); put16(out, 0); put16(out, 0); put16(out, nm); put16(out, nm); put32(out, (uint32_t)central.size());
  put32(out, cd_off); put16(out, 0);
  FILE* f = fopen(path, "wb");
  if (!f) return false;
  const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
  return (fclose(f) == 0) && ok;
}

}  // namespace cnmf

extern "C" int cnmf_b200_write_npz_batch(int64_t n_files, const char* const* paths, const void* const* data,
                                         const int64_t* data_bytes, const char* const* data_hdr,
                                         const int64_t* data_hdr_len, const char* const* index_npy,
                                         const int64_t* index_npy_len, const char* columns_comp, int64_t columns_comp_len,
                                         int64_t columns_raw_len, uint32_t columns_crc, int32_t level, int32_t threads) {
  using namespace cnmf;
  CNMF_NVTX("cnmf_b200_write_npz_batch");
  CNMF_CHECK_ARG(n_files >= 0 && paths && data && data_bytes && data_hdr && data_hdr_len && index_npy && index_npy_len &&
                     columns_comp,
                 "write_npz_batch: null pointer");
  for (int64_t i = 0; i < n_files; ++i)
    CNMF_CHECK_ARG(data_bytes[i] >= 0 && data_bytes[i] + data_hdr_len[i] < (1ll << 31) && index_npy_len[i] < (1ll << 31),
                   "write_npz_batch: a member of 2 GiB or more needs zip64 (file %lld)", (long long)i);
  int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  if (nt > 16) nt = 16;
  if (nt > n_files) nt = (int)(n_files > 0 ? n_files : 1);
  std::vector<int> failed(nt, 0);
  std::vector<std::thread> th;
  for (int w = 0; w < nt; ++w) {
    th.emplace_back([&, w] {
      for (int64_t i = w; i < n_files; i += nt) {
        ZipMember m[3];
        m[0].name = "data.npy"; m[0].shared = nullptr; m[0].shared_len = 0;
        m[0].raw_len = (uint32_t)(data_hdr_len[i] + data_bytes[i]);
        bool ok = deflate_raw(reinterpret_cast<const unsigned char*>(data_hdr[i]), (size_t)data_hdr_len[i],
                              static_cast<const unsigned char*>(data[i]), (size_t)data_bytes[i], level, m[0].comp, m[0].crc);
        m[1].name = "index.npy"; m[1].shared = nullptr; m[1].shared_len = 0; m[1].raw_len = (uint32_t)index_npy_len[i];
        ok = ok && deflate_raw(reinterpret_cast<const unsigned char*>(index_npy[i]), (size_t)index_npy_len[i], nullptr, 0,
                               level, m[1].comp, m[1].crc);
        m[2].name = "columns.npy"; m[2].shared = columns_comp; m[2].shared_len = (uint32_t)columns_comp_len;
        m[2].raw_len = (uint32_t)columns_raw_len; m[2].crc = columns_crc;
        ok = ok && write_zip_file(paths[i], m, 3);
        if (!ok) failed[w] = 1;
      }
    });
  }
  for (auto& x : th) x.join();
  for (int w = 0; w < nt; ++w)
    if (failed[w]) { set_error("write_npz_batch: compressing or writing a file failed"); return CNMF_B200_ERR_INVALID; }
  return 0;
}

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
