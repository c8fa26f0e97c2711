// links_b200/csrc/host/seq_stream.hpp -- streaming FASTA/FASTQ reader for the LINKS_CPP host.
//
// Hands the C-ABI what btllib::SeqReader hands LINKS in LONG_MODE (reference
// src/LINKS.hpp:404-468): records in file order, sequence bytes untouched (no case folding),
// multi-line FASTA joined, 0-based ordinal per file.  Compressed inputs are piped through the
// same external tools btllib forks (reference btllib/include/btllib/data_stream.hpp:36-44,112-118)
// keyed on the file suffix.  Records are delivered in BATCHES laid out exactly as
// links_reads_push / links_contigs_push take them: one concatenated ASCII buffer + n+1 offsets.
#pragma once
#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include <sys/wait.h>
#include <unistd.h>

namespace links_host {

struct SeqBatch {
  char* seq = nullptr;          // caller-provided buffer (pinned if it came from links_host_alloc)
  size_t cap = 0;               // bytes available in seq
  std::vector<uint64_t> off;    // n+1 offsets into seq
  uint64_t first_ordinal = 0;   // ordinal of record 0 of this batch
  // optional: called when ONE record needs more than cap bytes (the batch then holds nothing but the first `keep` bytes
  // of that record at seq[0..keep)); it may point seq at a larger buffer holding the same bytes and raise cap
  std::function<bool(SeqBatch&, size_t need, size_t keep)> grow;
  size_t n() const { return off.empty() ? 0 : off.size() - 1; }
  size_t bytes() const { return off.empty() ? 0 : size_t(off.back()); }
  void clear(uint64_t next_ordinal)
  {
    off.assign(1, 0);
    first_ordinal = next_ordinal;
  }
};

inline bool ends_with(const std::string& s, const char* suf)
{
  const size_t n = strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

inline std::string shell_quote(const std::string& s)
{
  std::string q = "'";
  for (char c : s) { if (c == '\'') q += "'\\''"; else q += c; }
  return q + "'";
}

// length of [p, p+l) without its trailing whitespace: btllib's readers rtrim() the sequence after every line they append
// (seq_reader_multiline_fasta_module.hpp:108,190,201; util.hpp:83-101: std::isspace), so blanks and tabs before the newline
// never become bases, while blanks inside a line do (as invalid ones)
inline size_t rstrip_len(const char* p, size_t l)
{
  while (l && std::isspace(static_cast<unsigned char>(p[l - 1]))) --l;
  return l;
}

// Opens path directly or through a decompressor pipe; closes the right way.
class ByteSource {
public:
  explicit ByteSource(const std::string& path)
  {
    const char* tool = nullptr;
    // pigz when the machine has it, else gzip -- btllib's own order of preference (data_stream.hpp:36-37)
    if (ends_with(path, ".gz") || ends_with(path, ".z")) tool = system("command -v pigz >/dev/null 2>&1") == 0 ? "pigz -dc" : "gzip -dc";
    else if (ends_with(path, ".bz2")) tool = "bzip2 -dc";
    else if (ends_with(path, ".xz")) tool = "xz -dc";
    else if (ends_with(path, ".zst")) tool = "zstd -dc";
    path_ = path;
    if (tool) {
      // popen() succeeds whatever the file is: check it ourselves, as btllib's DataSource does before it forks
      // (data_stream.hpp:112-118), and check the decompressor's exit status in finish()
      if (access(path.c_str(), R_OK) != 0) throw std::runtime_error("cannot open " + path);
      const std::string cmd = std::string(tool) + " " + shell_quote(path);
      f_ = popen(cmd.c_str(), "r");
      piped_ = true;
    } else {
      f_ = fopen(path.c_str(), "rb");
    }
    if (!f_) throw std::runtime_error("cannot open " + path);
  }
  ~ByteSource()
  {
    if (!f_) return;
    if (piped_) pclose(f_); else fclose(f_);
  }
  size_t read(char* dst, size_t n) { return fread(dst, 1, n, f_); }
  // after the last read: a decompressor that did not exit with status 0 (truncated or corrupt archive, missing tool)
  // has delivered partial data -- an error, never a shorter input
  void finish()
  {
    if (!f_) return;
    FILE* f = f_;
    f_ = nullptr;
    if (!piped_) { fclose(f); return; }
    const int st = pclose(f);
    if (st == -1 || !WIFEXITED(st) || WEXITSTATUS(st) != 0)
      throw std::runtime_error("decompressing " + path_ + " failed (exit status " + std::to_string(st == -1 ? -1 : WIFEXITED(st) ? WEXITSTATUS(st) : 128) + ")");
  }
  ByteSource(const ByteSource&) = delete;
  ByteSource& operator=(const ByteSource&) = delete;

private:
  FILE* f_ = nullptr;
  bool piped_ = false;
  std::string path_;
};

// Parses one file, appending records to `batch`; calls flush(batch) whenever the batch is full
// (by bytes or by record count) and once more at the end of the file if `flush_at_end`.
// A single record larger than the whole batch buffer is an error (the caller sizes the buffer).
// Returns the number of records read from this file.
inline uint64_t stream_file(const std::string& path, SeqBatch& batch, size_t max_records, bool flush_at_end,
                            const std::function<void(SeqBatch&)>& flush)
{
  ByteSource src(path);
  std::vector<char> buf(8u << 20);
  enum State { START, FA_HEADER, FA_SEQ, FQ_HEADER, FQ_SEQ, FQ_PLUS, FQ_QUAL } st = START;
  uint64_t n_records = 0;
  uint64_t cur_len = 0;    // bases of the open record already copied into batch.seq
  uint64_t qual_left = 0;  // FASTQ: quality characters still to skip
  bool open_rec = false;
  bool line_start = true;
  std::string pend_ws;     // trailing whitespace of a sequence line cut by the read buffer: bases only if the line goes on

  auto ordinal_next = [&]() { return batch.first_ordinal + batch.n(); };
  auto close_record = [&]() {
    batch.off.push_back(batch.off.back() + cur_len);
    ++n_records;
    open_rec = false;
    cur_len = 0;
    if (batch.n() >= max_records) { flush(batch); batch.clear(ordinal_next()); }
  };
  auto append = [&](const char* p, size_t n) {
    if (batch.bytes() + cur_len + n > batch.cap) {
      if (batch.n() > 0) {  // move the open record to the front of a fresh batch
        const uint64_t start = batch.off.back();
        flush(batch);
        memmove(batch.seq, batch.seq + start, cur_len);
        batch.clear(ordinal_next());
      }
      if (cur_len + n > batch.cap && (!batch.grow || !batch.grow(batch, cur_len + n, cur_len) || cur_len + n > batch.cap))
        throw std::runtime_error("record larger than the staging buffer in " + path);
    }
    memcpy(batch.seq + batch.off.back() + cur_len, p, n);
    cur_len += n;
  };

  for (;;) {
    const size_t got = src.read(buf.data(), buf.size());
    if (!got) break;
    size_t i = 0;
    while (i < got) {
      const char* nl = static_cast<const char*>(memchr(buf.data() + i, '\n', got - i));
      const size_t end = nl ? size_t(nl - buf.data()) : got;  // [i, end) is (part of) a line
      size_t len = end - i;
      const char* p = buf.data() + i;
      if (line_start && len) {
        // first byte of a line decides the state transition
        if (st == START || st == FA_SEQ) {
          if (*p == '>') { if (open_rec) close_record(); st = FA_HEADER; }
          else if (*p == '@' && st == START) st = FQ_HEADER;
          else if (st == START) throw std::runtime_error("unrecognised sequence format in " + path);
        } else if (st == FQ_SEQ) {
          // FASTQ sequence lines may wrap; '+' at line start ends the sequence
          if (*p == '+') st = FQ_PLUS;
        }
      }
      switch (st) {
      case FA_SEQ:
      case FQ_SEQ: {
        const size_t l = rstrip_len(p, len);
        if (l) {
          open_rec = true;
          if (!pend_ws.empty()) { append(pend_ws.data(), pend_ws.size()); pend_ws.clear(); }
          append(p, l);
        }
        if (nl) pend_ws.clear();                  // whitespace before the newline: dropped
        else pend_ws.append(p + l, len - l);      // the line continues in the next buffer
        break;
      }
      case FQ_QUAL: {
        size_t l = len;
        if (nl) l = rstrip_len(p, l);
        qual_left = l >= qual_left ? 0 : qual_left - l;
        break;
      }
      default: break;  // header / plus lines are skipped
      }
      if (!nl) { line_start = false; break; }
      // end of line: advance the state machine
      line_start = true;
      i = end + 1;
      switch (st) {
      case FA_HEADER: st = FA_SEQ; open_rec = true; cur_len = 0; break;
      case FQ_HEADER: st = FQ_SEQ; open_rec = true; cur_len = 0; break;
      case FQ_PLUS: st = FQ_QUAL; qual_left = cur_len; if (!qual_left) { close_record(); st = START; } break;
      case FQ_QUAL: if (!qual_left) { close_record(); st = START; } break;
      default: break;
      }
    }
  }
  src.finish();
  if (open_rec && (st == FA_SEQ || st == FA_HEADER || st == FQ_SEQ || st == FQ_PLUS || st == FQ_QUAL)) close_record();
  if (flush_at_end && batch.n()) { flush(batch); batch.clear(ordinal_next()); }
  return n_records;
}

}  // namespace links_host
