// Multi-threaded CSV writer for the per-epoch synthetic table (host code).
// The reference writes 40 000 x 42 cells with pandas inside the timed round
// (Server/dtds/distributed.py:586-590).  At 8 clients a federated round takes ~4 ms, so
// the federator has to emit ~270 CSVs per second: the writer is built for CPU time per file.
//  * every thread formats a row range straight into its slice of one worst-case-sized buffer
//    (pointer bumps, no per-cell capacity checks or allocations);
//  * the buffer belongs to the calling thread and is reused by its next call (the federator's
//    writer pool has a few long-lived threads), so steady state takes no page faults;
//  * category strings are quoted and measured once per call, a cell is then one memcpy;
//  * floats use std::to_chars (shortest string that round-trips, like Python's repr), with an
//    integer fast path for integral values (counts, ages, ... written as "123.0").
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <deque>
#include <string>
#include <thread>
#include <vector>

namespace {

struct StrEntry { const char* p; uint32_t len; };

inline bool needs_quote(const char* s) {
  for (const char* p = s; *p; ++p)
    if (*p == ',' || *p == '"' || *p == '\n' || *p == '\r') return true;
  return false;
}

void quote_into(std::string& out, const char* s) {
  if (!needs_quote(s)) { out.append(s); return; }
  out.push_back('"');
  for (const char* p = s; *p; ++p) { if (*p == '"') out.push_back('"'); out.push_back(*p); }
  out.push_back('"');
}

// integral |v| < 2^24 (float) / 2^53 (double): digits + ".0"; returns nullptr when not applicable
template <typename T>
inline char* put_integral(char* o, T v, T limit) {
  if (!(v > -limit && v < limit)) return nullptr;
  const long long i = (long long)v;
  if ((T)i != v) return nullptr;
  if (i == 0 && std::signbit(v)) { std::memcpy(o, "-0.0", 4); return o + 4; }
  auto res = std::to_chars(o, o + 24, i);
  res.ptr[0] = '.'; res.ptr[1] = '0';
  return res.ptr + 2;
}

template <typename T>
inline char* put_real(char* o, T v, T int_limit) {
  if (std::isnan(v)) return o;                               // pandas writes NaN as an empty cell
  if (std::isinf(v)) { const char* s = v > 0 ? "inf" : "-inf"; size_t n = std::strlen(s); std::memcpy(o, s, n); return o + n; }
  if (char* e = put_integral<T>(o, v, int_limit)) return e;
  auto res = std::to_chars(o, o + 40, v);
  bool plain = true;
  for (char* p = o; p < res.ptr; ++p)
    if (*p == '.' || *p == 'e') { plain = false; break; }
  if (plain) { res.ptr[0] = '.'; res.ptr[1] = '0'; return res.ptr + 2; }   // keep float-ness (repr style)
  return res.ptr;
}

}  // namespace

extern "C" {

// kinds: 0 = float64 column, 1 = float32 column, 2 = int64 column, 3 = int32 codes into
// str_tables[c] (codes index the table directly; its length is taken from the largest code),
// 4 = float32 column holding category CODES (the generator's numeric table): code = clamp((int)v, 0, n_classes[c]-1).
// `strides` (elements between consecutive rows of a column; NULL = 1 everywhere) lets kinds 1 and 4 read the
// columns of a row-major [n_rows, ld] float32 table in place -- no per-column host copies / dtype passes.
// A negative `n_threads` means |n_threads| threads and NO header line (the file is one part of a table whose parts
// are concatenated by the caller).
int ftg_csv_write_strided(const char* path, int n_rows, int n_cols, const char* const* names, const int* kinds,
                          const void* const* columns, const int* strides, const int* n_classes,
                          const char* const* const* str_tables, int n_threads) {
  const bool with_header = n_threads >= 0;
  if (n_threads < 0) n_threads = -n_threads;
  if (n_rows < 0 || n_cols <= 0) return -2;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  if (n_threads > n_rows) n_threads = n_rows > 0 ? n_rows : 1;

  // ---- per-column preparation: quoted category strings, worst-case cell width
  std::vector<std::vector<StrEntry>> tables(n_cols);
  std::deque<std::string> quoted_store;                      // owns the strings that needed quoting (stable addresses)
  size_t row_bytes = 1;                                      // '\n'
  for (int c = 0; c < n_cols; ++c) {
    size_t w = 0;
    switch (kinds[c]) {
      case 0: w = 26; break;                                 // "-2.2250738585072014e-308" is 24 characters
      case 1: w = 16; break;                                 // "-1.17549435e-38" is 15
      case 2: w = 21; break;
      default: {
        int32_t mx = -1;
        if (kinds[c] == 4) {
          if (!n_classes || n_classes[c] < 1) return -5;
          mx = n_classes[c] - 1;
        } else {
          const int32_t* codes = static_cast<const int32_t*>(columns[c]);
          for (int r = 0; r < n_rows; ++r) {
            if (codes[r] < 0) return -3;
            if (codes[r] > mx) mx = codes[r];
          }
        }
        tables[c].resize((size_t)(mx + 1));
        for (int32_t k = 0; k <= mx; ++k) {
          const char* s = str_tables[c][k];
          if (needs_quote(s)) {
            quoted_store.emplace_back();
            quote_into(quoted_store.back(), s);
            tables[c][k] = {quoted_store.back().data(), (uint32_t)quoted_store.back().size()};
          } else {
            tables[c][k] = {s, (uint32_t)std::strlen(s)};
          }
          if (tables[c][k].len > w) w = tables[c][k].len;
        }
        break;
      }
    }
    row_bytes += w + 1;                                      // cell + ','
  }

  // ---- one worst-case-sized buffer per calling thread, reused across calls
  static thread_local std::vector<char> buffer;
  const size_t rows_per = ((size_t)n_rows + n_threads - 1) / n_threads;
  const size_t slice = rows_per * row_bytes;
  if (buffer.size() < slice * n_threads + 64) buffer.resize(slice * n_threads + 64);   // +64: to_chars end bounds stay inside
  char* const base = buffer.data();
  std::vector<size_t> used(n_threads, 0);

  auto work = [&](int t) {
    const long long r0 = (long long)rows_per * t;
    const long long r1 = std::min<long long>(n_rows, r0 + (long long)rows_per);
    char* o = base + slice * t;
    for (long long r = r0; r < r1; ++r) {
      for (int c = 0; c < n_cols; ++c) {
        if (c) *o++ = ',';
        switch (kinds[c]) {
          case 0: o = put_real<double>(o, static_cast<const double*>(columns[c])[r], 9007199254740992.0); break;
          case 1: o = put_real<float>(o, static_cast<const float*>(columns[c])[r * (long long)(strides ? strides[c] : 1)], 16777216.0f); break;
          case 4: {
            const float v = static_cast<const float*>(columns[c])[r * (long long)(strides ? strides[c] : 1)];
            // clamp in float space first: (int) of a huge / infinite float is undefined.  NaN / negative -> 0
            const float top = (float)(n_classes[c] - 1);
            const int code = (v >= top) ? n_classes[c] - 1 : (v > 0.f ? (int)v : 0);
            const StrEntry& e = tables[c][code];
            std::memcpy(o, e.p, e.len);
            o += e.len;
            break;
          }
          case 2: o = std::to_chars(o, o + 24, static_cast<const int64_t*>(columns[c])[r]).ptr; break;
          default: {
            const StrEntry& e = tables[c][static_cast<const int32_t*>(columns[c])[r]];
            std::memcpy(o, e.p, e.len);
            o += e.len;
            break;
          }
        }
      }
      *o++ = '\n';
    }
    used[t] = (size_t)(o - (base + slice * t));
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < n_threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();

  FILE* f = std::fopen(path, "wb");
  if (!f) return -1;
  std::string header;
  for (int c = 0; c < n_cols; ++c) { if (c) header.push_back(','); quote_into(header, names[c]); }
  header.push_back('\n');
  bool ok = !with_header || std::fwrite(header.data(), 1, header.size(), f) == header.size();
  for (int t = 0; t < n_threads && ok; ++t) ok = std::fwrite(base + slice * t, 1, used[t], f) == used[t];
  ok = (std::fclose(f) == 0) && ok;
  if (buffer.size() > ((size_t)256 << 20)) std::vector<char>().swap(buffer);   // do not pin huge tables
  return ok ? 0 : -4;
}

int ftg_csv_write(const char* path, int n_rows, int n_cols, const char* const* names, const int* kinds,
                  const void* const* columns, const char* const* const* str_tables, int n_threads) {
  return ftg_csv_write_strided(path, n_rows, n_cols, names, kinds, columns, nullptr, nullptr, str_tables, n_threads);
}

}  // extern "C"
