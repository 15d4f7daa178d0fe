// Host-side text -> token-row packer for the real-data path (utils/data.py: TextFileDataset).
//
// The reference tokenises wikitext with HF `tokenizers` (Rust) and pads every row to max_length
// (util.py:47-51).  Offline there are no vocabulary files, so rows are built with a deterministic
// word-hash vocabulary instead: bytes are ASCII-lower-cased, a word is a run of [a-z0-9_] or
// bytes >= 0x80 (UTF-8 continuation/lead bytes stay inside words), any other non-space byte is a
// token of its own; token id = first_id + FNV-1a(word) % (vocab - first_id).  One non-empty line
// per row, truncated to max_len, padded with pad_id.  Lines are split across threads.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <string>
#include <thread>
#include <vector>

namespace {

inline bool is_space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); }
inline bool is_word(unsigned char c) {
  return (c >= 'a' && c <= 'z') || (c >= '0' && c <= '9') || c == '_' || c >= 0x80;
}
inline unsigned char lower(unsigned char c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }

int encode_line(const std::string& line, long long vocab, long long first_id, int max_len, long long* out) {
  const unsigned long long span = static_cast<unsigned long long>(vocab - first_id);
  int n = 0;
  size_t i = 0;
  const size_t len = line.size();
  while (i < len && n < max_len) {
    unsigned char c = lower(static_cast<unsigned char>(line[i]));
    if (is_space(c)) { ++i; continue; }
    unsigned long long h = 0xcbf29ce484222325ull;
    if (is_word(c)) {
      while (i < len) {
        c = lower(static_cast<unsigned char>(line[i]));
        if (!is_word(c)) break;
        h = (h ^ c) * 0x100000001b3ull;
        ++i;
      }
    } else {
      h = (h ^ c) * 0x100000001b3ull;
      ++i;
    }
    out[n++] = first_id + static_cast<long long>(h % span);
  }
  return n;
}

}  // namespace

// Fills ids[rows, max_len] (pad_id beyond each row's length) and lengths[rows] from the non-empty
// lines of `path`.  Returns the number of rows written (<= max_rows), or -1 if the file cannot be
// read, -2 on bad arguments.
extern "C" long long dtds_text_rows(const char* path, long long vocab, long long first_id, int max_len,
                                    long long pad_id, long long* ids, int* lengths, long long max_rows,
                                    int threads) {
  if (path == nullptr || ids == nullptr || lengths == nullptr || max_len <= 0 || vocab <= first_id || first_id < 0)
    return -2;
  std::ifstream f(path, std::ios::binary);
  if (!f) return -1;
  std::vector<std::string> lines;
  std::string line;
  while (static_cast<long long>(lines.size()) < max_rows && std::getline(f, line)) {
    bool blank = true;
    for (unsigned char c : line)
      if (!is_space(c)) { blank = false; break; }
    if (!blank) lines.push_back(std::move(line));
  }
  const long long rows = static_cast<long long>(lines.size());
  if (threads <= 0) threads = static_cast<int>(std::min<long long>(std::max(1u, std::thread::hardware_concurrency()), 16));
  threads = static_cast<int>(std::max<long long>(1, std::min<long long>(threads, (rows + 255) / 256)));
  auto work = [&](long long lo, long long hi) {
    for (long long r = lo; r < hi; ++r) {
      long long* row = ids + r * max_len;
      const int n = encode_line(lines[r], vocab, first_id, max_len, row);
      for (int k = n; k < max_len; ++k) row[k] = pad_id;
      lengths[r] = n;
    }
  };
  std::vector<std::thread> pool;
  const long long per = (rows + threads - 1) / threads;
  for (int t = 0; t < threads; ++t) {
    const long long lo = t * per, hi = std::min(rows, lo + per);
    if (lo < hi) pool.emplace_back(work, lo, hi);
  }
  for (auto& th : pool) th.join();
  return rows;
}
