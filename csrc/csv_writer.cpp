// Multi-threaded CSV writer for the per-epoch synthetic table (host code).
// The reference writes 40 000 x 42 cells with pandas inside the timed round
// (Server/dtds/distributed.py:586-590); here each thread formats a row range
// with std::to_chars (shortest round-trip, like Python's repr) into its own buffer.
#include <charconv>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <cstdint>
#include <cmath>

extern "C" {

// kinds: 0 = float64 column, 1 = float32 column, 2 = int64 column, 3 = int32 codes into
// str_tables[c] (NULL-terminated array of C strings is not required; codes index directly)
int ftg_csv_write(const char* path, int n_rows, int n_cols, const char* const* names, const int* kinds,
                  const void* const* columns, const char* const* const* str_tables, int n_threads) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  std::vector<std::string> bufs(n_threads);
  auto needs_quote = [](const char* s) {
    for (const char* p = s; *p; ++p)
      if (*p == ',' || *p == '"' || *p == '\n' || *p == '\r') return true;
    return false;
  };
  auto put_str = [&](std::string& out, const char* s) {
    if (!needs_quote(s)) { out.append(s); return; }
    out.push_back('"');
    for (const char* p = s; *p; ++p) { if (*p == '"') out.push_back('"'); out.push_back(*p); }
    out.push_back('"');
  };
  auto put_float = [](std::string& out, float v) {
    if (std::isnan(v)) return;
    if (std::isinf(v)) { out.append(v > 0 ? "inf" : "-inf"); return; }
    char tmp[24];
    auto res = std::to_chars(tmp, tmp + sizeof(tmp), v);      // shortest string that round-trips the float32
    bool plain = true;
    for (char* p = tmp; p < res.ptr; ++p)
      if (*p == '.' || *p == 'e') { plain = false; break; }
    out.append(tmp, res.ptr - tmp);
    if (plain) out.append(".0");
  };
  auto put_double = [](std::string& out, double v) {
    if (std::isnan(v)) return;                        // pandas writes NaN as empty
    if (std::isinf(v)) { out.append(v > 0 ? "inf" : "-inf"); return; }
    char tmp[40];
    auto res = std::to_chars(tmp, tmp + sizeof(tmp), v);
    bool plain = true;
    for (char* p = tmp; p < res.ptr; ++p)
      if (*p == '.' || *p == 'e' || *p == 'n' || *p == 'i') { plain = false; break; }
    out.append(tmp, res.ptr - tmp);
    if (plain) out.append(".0");                      // keep float-ness (repr style)
  };
  auto work = [&](int t) {
    long long r0 = (long long)n_rows * t / n_threads, r1 = (long long)n_rows * (t + 1) / n_threads;
    std::string& out = bufs[t];
    out.reserve((size_t)(r1 - r0) * n_cols * 14);
    for (long long r = r0; r < r1; ++r) {
      for (int c = 0; c < n_cols; ++c) {
        if (c) out.push_back(',');
        switch (kinds[c]) {
          case 0: put_double(out, static_cast<const double*>(columns[c])[r]); break;
          case 1: put_float(out, static_cast<const float*>(columns[c])[r]); break;
          case 2: {
            char tmp[24];
            auto res = std::to_chars(tmp, tmp + sizeof(tmp), static_cast<const int64_t*>(columns[c])[r]);
            out.append(tmp, res.ptr - tmp);
            break;
          }
          default: put_str(out, str_tables[c][static_cast<const int32_t*>(columns[c])[r]]); break;
        }
      }
      out.push_back('\n');
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < n_threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  FILE* f = std::fopen(path, "wb");
  if (!f) return -1;
  std::string header;
  for (int c = 0; c < n_cols; ++c) { if (c) header.push_back(','); put_str(header, names[c]); }
  header.push_back('\n');
  std::fwrite(header.data(), 1, header.size(), f);
  for (auto& b : bufs) std::fwrite(b.data(), 1, b.size(), f);
  std::fclose(f);
  return 0;
}

}  // extern "C"
