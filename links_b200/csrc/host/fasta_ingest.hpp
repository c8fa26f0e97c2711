// links_b200/csrc/host/fasta_ingest.hpp -- multi-threaded ingest of plain (uncompressed) FASTA and four-line FASTQ files.
//
// SURVEY 8(f) row 1.  The reference reads sequences with btllib::SeqReader worker threads feeding LINKS'
// InputWorker (reference src/LINKS.hpp:404-468); with the hot path on the GPU the single-threaded line
// scanner of seq_stream.hpp becomes the slowest stage of LINKS_CPP (about 1.5 GB/s against a device that
// takes tens of GB/s).  This reader maps the file, indexes a window of it with T threads (record
// boundaries and base counts), copies the sequence lines with T threads straight into one of two pinned
// staging buffers at their final offsets, and hands finished batches to a single pusher thread, so that
// parsing batch i+1 overlaps the host-to-device copy and the kernels of batch i.
//
// Record semantics are those of seq_stream.hpp (and so of SeqReader in LONG_MODE): records in file order,
// bases untouched, lines joined, trailing '\r' dropped, a header directly followed by another header is an
// empty record.  FASTQ is taken when the WHOLE file is strict four-line FASTQ (one sequence line, one quality line of
// the same length per record, no blank lines but at the end): the constructor indexes and verifies it with T threads --
// every thread finds its first record by "a line starting with '@' whose second successor starts with '+'", and the
// pieces must meet exactly -- and anything else (wrapped FASTQ, truncated records, mixed files) is left to the serial
// reader, whose behaviour is the definition.  Compressed or non-regular inputs are not handled here either: usable()
// is false and the caller falls back to stream_file().
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "seq_stream.hpp"  // rstrip_len: the line rule both readers share

namespace links_host {

class ParallelFastaIngest {
public:
  // sink(seq, off, n, first_ordinal): one batch laid out as links_reads_push / links_contigs_push take it.
  // Called from ONE thread, in file order; the buffer may be reused as soon as it returns.
  using Sink = std::function<void(const char*, const uint64_t*, size_t, uint64_t)>;

  // min_segment: least bytes of file per indexing thread (tests shrink it so that small files are cut into many pieces)
  ParallelFastaIngest(const std::string& path, unsigned threads, uint64_t min_segment = 1u << 16)
    : path_(path), threads_(std::max(threads, 1u)), min_segment_(std::max<uint64_t>(min_segment, 1))
  {
    for (const char* suf : {".gz", ".bz2", ".xz", ".zst", ".zip", ".lrz", ".7z"})
      if (path.size() >= strlen(suf) && path.compare(path.size() - strlen(suf), strlen(suf), suf) == 0) return;
    fd_ = open(path.c_str(), O_RDONLY);
    if (fd_ < 0) return;
    struct stat st;
    if (fstat(fd_, &st) != 0 || !S_ISREG(st.st_mode) || st.st_size == 0) return;
    size_ = uint64_t(st.st_size);
    void* m = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (m == MAP_FAILED) return;
    data_ = static_cast<const char*>(m);
    madvise(const_cast<char*>(data_), size_, MADV_SEQUENTIAL);
    // first non-blank line decides the format, as in stream_file()
    uint64_t p = 0;
    while (p < size_ && data_[p] == '\n') ++p;  // (a line holding only '\r' is not blank to stream_file(): it rejects the file)
    if (p < size_ && data_[p] == '>' && (p == 0 || data_[p - 1] == '\n')) { first_header_ = p; usable_ = true; }
    else if (p < size_ && data_[p] == '@' && (p == 0 || data_[p - 1] == '\n')) { first_header_ = p; fastq_ = true; usable_ = index_fastq(); }
  }
  ~ParallelFastaIngest()
  {
    if (data_) munmap(const_cast<char*>(data_), size_);
    if (fd_ >= 0) close(fd_);
  }
  ParallelFastaIngest(const ParallelFastaIngest&) = delete;
  ParallelFastaIngest& operator=(const ParallelFastaIngest&) = delete;

  bool usable() const { return usable_; }
  bool is_fastq() const { return fastq_; }
  void set_window_bytes(uint64_t w) { window_ = std::max<uint64_t>(w, 1); }  // tests shrink it to exercise the seams

  // Streams every record to `sink` in batches of at most cap bytes / max_records records, alternating between
  // buf[0] and buf[1].  Returns the number of records.  A record larger than cap: `grow(need)` is called with no batch
  // in flight and may replace buf[0], buf[1] and cap by larger buffers (returning true); without it, or if it returns
  // false, std::runtime_error is thrown.
  using Grow = std::function<bool(size_t need, char* buf[2], size_t& cap)>;
  uint64_t run(char* buf[2], size_t& cap, size_t max_records, uint64_t first_ordinal, const Sink& sink, const Grow& grow = nullptr);

private:
  struct Rec { uint64_t seq_begin, end, len; };  // file span of the sequence lines, bases in them

  uint64_t line_start_at_or_after(uint64_t s) const
  {
    if (s == 0 || data_[s - 1] == '\n') return s;
    const void* nl = memchr(data_ + s, '\n', size_ - s);
    return nl ? uint64_t(static_cast<const char*>(nl) - data_) + 1 : size_;
  }
  // records whose header starts in [s, e); s is a line start
  void index_segment(uint64_t s, uint64_t e, std::vector<Rec>& out) const;
  void copy_record(const Rec& r, char* dst) const;
  template <class F> void parallel(unsigned n, F&& f) const
  {
    if (n <= 1) { f(0u); return; }
    std::vector<std::thread> th;
    th.reserve(n - 1);
    for (unsigned t = 1; t < n; ++t) th.emplace_back([&f, t] { f(t); });
    f(0u);
    for (auto& x : th) x.join();
  }

  std::string path_;
  unsigned threads_;
  uint64_t min_segment_;
  int fd_ = -1;
  const char* data_ = nullptr;
  uint64_t size_ = 0, first_header_ = 0, window_ = 256ull << 20;
  bool usable_ = false, fastq_ = false;
  std::vector<Rec> fq_recs_;  // FASTQ: every record of the file (index_fastq)

  // end of the line starting at p (position of its '\n', or size_ when the last line has none)
  uint64_t line_end(uint64_t p) const
  {
    const void* nl = memchr(data_ + p, '\n', size_ - p);
    return nl ? uint64_t(static_cast<const char*>(nl) - data_) : size_;
  }
  bool rest_is_blank(uint64_t p) const
  {
    for (; p < size_; ++p)
      if (data_[p] != '\n') return false;  // a "\r\n" blank line is an error to stream_file(): let it say so
    return true;
  }
  uint64_t fastq_record_at_or_after(uint64_t s) const;
  bool index_fastq_segment(uint64_t p, uint64_t e, std::vector<Rec>& out, uint64_t& end_pos) const;
  bool index_fastq();
};

inline void ParallelFastaIngest::index_segment(uint64_t s, uint64_t e, std::vector<Rec>& out) const
{
  uint64_t p = s;
  while (p < e && data_[p] != '>') {  // skip the tail of a record that belongs to the previous segment
    const void* nl = memchr(data_ + p, '\n', size_ - p);
    p = nl ? uint64_t(static_cast<const char*>(nl) - data_) + 1 : size_;
  }
  while (p < e) {
    const void* hnl = memchr(data_ + p, '\n', size_ - p);
    if (!hnl) break;  // header line without a newline at end of file: no record (as stream_file)
    uint64_t q = uint64_t(static_cast<const char*>(hnl) - data_) + 1;
    Rec r{q, q, 0};
    while (q < size_ && data_[q] != '>') {
      const void* nl = memchr(data_ + q, '\n', size_ - q);
      const uint64_t lend = nl ? uint64_t(static_cast<const char*>(nl) - data_) : size_;
      const uint64_t l = links_host::rstrip_len(data_ + q, lend - q);
      r.len += l;
      q = nl ? lend + 1 : size_;
    }
    r.end = q;
    out.push_back(r);
    p = q;
  }
}

// first position >= s that starts a line with '@' whose second successor line starts with '+' (size_ if none).
// In four-line FASTQ a quality line may start with '@' too, but two lines below it is a sequence line, never a '+'.
inline uint64_t ParallelFastaIngest::fastq_record_at_or_after(uint64_t s) const
{
  uint64_t p = line_start_at_or_after(s);
  while (p < size_) {
    const uint64_t e0 = line_end(p);
    if (data_[p] == '@' && e0 < size_) {
      const uint64_t e1 = line_end(e0 + 1);
      if (e1 < size_ && e1 + 1 < size_ && data_[e1 + 1] == '+') return p;
    }
    if (e0 >= size_) break;
    p = e0 + 1;
  }
  return size_;
}

// records starting in [p, e), p being a record start; false = not strict four-line FASTQ here
inline bool ParallelFastaIngest::index_fastq_segment(uint64_t p, uint64_t e, std::vector<Rec>& out, uint64_t& end_pos) const
{
  auto stripped = [&](uint64_t b, uint64_t le, bool has_nl) -> uint64_t {  // line length without the '\r' of a CRLF
    (void)has_nl;
    return links_host::rstrip_len(data_ + b, le - b);
  };
  while (p < e && p < size_) {
    if (data_[p] != '@') {  // only blank lines may follow the last record
      if (!rest_is_blank(p)) return false;
      p = size_;
      break;
    }
    const uint64_t e0 = line_end(p);
    if (e0 >= size_) return false;
    const uint64_t b1 = e0 + 1, e1 = line_end(b1);
    if (e1 >= size_ || (b1 < e1 && (data_[b1] == '+' || data_[b1] == '@' || data_[b1] == '>'))) return false;
    const uint64_t b2 = e1 + 1;
    if (b2 >= size_ || data_[b2] != '+') return false;
    const uint64_t e2 = line_end(b2);
    if (e2 >= size_) return false;
    const uint64_t b3 = e2 + 1, e3 = line_end(b3);
    const uint64_t seq_len = stripped(b1, e1, true);
    if (seq_len == 0) return false;  // the serial reader closes such a record at the '+' line: leave it to that reader
    if (stripped(b3, e3, e3 < size_) != seq_len) return false;  // wrapped or truncated quality
    out.push_back(Rec{b1, e1, seq_len});
    p = e3 < size_ ? e3 + 1 : size_;
  }
  end_pos = p;
  return true;
}

inline bool ParallelFastaIngest::index_fastq()
{
  const uint64_t span = size_ - first_header_;
  const unsigned T = unsigned(std::min<uint64_t>(threads_, std::max<uint64_t>(span / min_segment_, 1)));
  std::vector<std::vector<Rec>> per(T);
  std::vector<uint64_t> first(T, 0), last(T, 0);
  std::vector<char> ok(T, 1);
  parallel(T, [&](unsigned t) {
    const uint64_t s = first_header_ + span * t / T, e = t + 1 == T ? size_ : first_header_ + span * (t + 1) / T;
    const uint64_t p = t == 0 ? first_header_ : fastq_record_at_or_after(s);
    first[t] = last[t] = p;
    if (p < e) ok[t] = index_fastq_segment(p, e, per[t], last[t]) ? 1 : 0;
  });
  for (unsigned t = 0; t < T; ++t) {
    if (!ok[t]) return false;
    if (t && first[t] != last[t - 1]) return false;  // the pieces must meet exactly, or the heuristic start was wrong
  }
  if (last[T - 1] != size_ && !rest_is_blank(last[T - 1])) return false;
  size_t n = 0;
  for (auto& v : per) n += v.size();
  fq_recs_.reserve(n);
  for (auto& v : per) fq_recs_.insert(fq_recs_.end(), v.begin(), v.end());
  return true;
}

inline void ParallelFastaIngest::copy_record(const Rec& r, char* dst) const
{
  uint64_t q = r.seq_begin;
  while (q < r.end) {
    const void* nl = memchr(data_ + q, '\n', r.end - q);
    const uint64_t lend = nl ? uint64_t(static_cast<const char*>(nl) - data_) : r.end;
    const uint64_t l = links_host::rstrip_len(data_ + q, lend - q);
    memcpy(dst, data_ + q, l);
    dst += l;
    q = nl ? lend + 1 : r.end;
  }
}

inline uint64_t ParallelFastaIngest::run(char* buf[2], size_t& cap, size_t max_records, uint64_t first_ordinal, const Sink& sink, const Grow& grow)
{
  if (!usable_) throw std::runtime_error("ParallelFastaIngest: " + path_ + " is not a plain FASTA file");
  // pusher thread: takes filled buffers in order
  struct Job { int b; std::vector<uint64_t> off; uint64_t first; };
  std::mutex mu;
  std::condition_variable cv;
  bool busy[2] = {false, false}, stop = false;
  std::vector<Job> queue;  // at most two entries
  std::string error;
  std::thread pusher([&] {
    for (;;) {
      Job j;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return stop || !queue.empty(); });
        if (queue.empty()) return;
        j = std::move(queue.front());
        queue.erase(queue.begin());
      }
      try {
        if (error.empty()) sink(buf[j.b], j.off.data(), j.off.size() - 1, j.first);
      } catch (const std::exception& e) {
        std::lock_guard<std::mutex> lk(mu);
        error = e.what();
      }
      {
        std::lock_guard<std::mutex> lk(mu);
        busy[j.b] = false;
      }
      cv.notify_all();
    }
  });
  auto finish = [&] {
    {
      std::lock_guard<std::mutex> lk(mu);
      stop = true;
    }
    cv.notify_all();
    pusher.join();
  };

  uint64_t n_records = 0, ordinal = first_ordinal;
  int next_buf = 0;
  try {
    uint64_t pos = first_header_;
    std::vector<std::vector<Rec>> per(threads_);
    std::vector<Rec> recs;
    while (pos < size_) {
      if (fastq_) {  // indexed and verified as a whole by the constructor
        recs.swap(fq_recs_);
        pos = size_;
        if (recs.empty()) break;
      } else {
        const uint64_t wend = std::min(size_, pos + window_);
        const unsigned T = unsigned(std::min<uint64_t>(threads_, std::max<uint64_t>((wend - pos) / min_segment_, 1)));
        parallel(T, [&](unsigned t) {
          per[t].clear();
          const uint64_t s = t == 0 ? pos : line_start_at_or_after(pos + (wend - pos) * t / T);
          const uint64_t e = t + 1 == T ? wend : line_start_at_or_after(pos + (wend - pos) * (t + 1) / T);
          if (s < e) index_segment(s, e, per[t]);
        });
        recs.clear();
        for (unsigned t = 0; t < T; ++t) recs.insert(recs.end(), per[t].begin(), per[t].end());
        if (recs.empty()) break;  // only an unterminated header line was left
        pos = recs.back().end;
      }
      // cut the window's records into batches and fill them
      size_t i0 = 0;
      while (i0 < recs.size()) {
        Job j;
        j.off.assign(1, 0);
        size_t i = i0;
        for (; i < recs.size() && i - i0 < max_records; ++i) {
          if (j.off.back() + recs[i].len > cap) break;
          j.off.push_back(j.off.back() + recs[i].len);
        }
        if (i == i0) {  // this record alone does not fit: wait for the pushes in flight, then ask for larger buffers
          {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return (!busy[0] && !busy[1] && queue.empty()) || !error.empty(); });
            if (!error.empty()) throw std::runtime_error(error);
          }
          if (!grow || !grow(size_t(recs[i0].len), buf, cap) || recs[i0].len > cap)
            throw std::runtime_error("record larger than the staging buffer in " + path_);
          continue;
        }
        j.b = next_buf;
        j.first = ordinal;
        {
          std::unique_lock<std::mutex> lk(mu);
          cv.wait(lk, [&] { return !busy[j.b] || !error.empty(); });
          if (!error.empty()) throw std::runtime_error(error);
          busy[j.b] = true;
        }
        const size_t n = i - i0;
        const uint64_t bytes = j.off.back();
        const unsigned TC = unsigned(std::min<uint64_t>(threads_, std::max<uint64_t>(bytes >> 16, 1)));
        char* dst = buf[j.b];
        const uint64_t* off = j.off.data();
        parallel(TC, [&](unsigned t) {
          // records [a, b) whose first byte falls into this thread's share of the batch
          const uint64_t lo = bytes * t / TC, hi = bytes * (t + 1) / TC;
          size_t a = size_t(std::lower_bound(off, off + n, lo) - off);
          size_t b = t + 1 == TC ? n : size_t(std::lower_bound(off, off + n, hi) - off);
          for (size_t r = a; r < b; ++r) copy_record(recs[i0 + r], dst + off[r]);
        });
        {
          std::lock_guard<std::mutex> lk(mu);
          queue.push_back(std::move(j));
        }
        cv.notify_all();
        ordinal += n;
        n_records += n;
        next_buf ^= 1;
        i0 = i;
      }
      // the index of the next window may only reuse `recs` once both in-flight batches are copied: they are (copy is synchronous)
    }
    {
      std::unique_lock<std::mutex> lk(mu);
      cv.wait(lk, [&] { return (!busy[0] && !busy[1] && queue.empty()) || !error.empty(); });
    }
  } catch (...) {
    finish();
    throw;
  }
  finish();
  if (!error.empty()) throw std::runtime_error(error);
  return n_records;
}

}  // namespace links_host
