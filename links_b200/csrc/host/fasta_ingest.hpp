// links_b200/csrc/host/fasta_ingest.hpp -- multi-threaded ingest of plain (uncompressed) FASTA files.
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
// empty record.  FASTQ, compressed or non-regular inputs are not handled here: usable() is false and the
// caller falls back to stream_file().
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

namespace links_host {

class ParallelFastaIngest {
public:
  // sink(seq, off, n, first_ordinal): one batch laid out as links_reads_push / links_contigs_push take it.
  // Called from ONE thread, in file order; the buffer may be reused as soon as it returns.
  using Sink = std::function<void(const char*, const uint64_t*, size_t, uint64_t)>;

  ParallelFastaIngest(const std::string& path, unsigned threads) : path_(path), threads_(std::max(threads, 1u))
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
    while (p < size_ && (data_[p] == '\n' || (data_[p] == '\r' && p + 1 < size_ && data_[p + 1] == '\n'))) ++p;
    if (p < size_ && data_[p] == '>' && (p == 0 || data_[p - 1] == '\n')) { first_header_ = p; usable_ = true; }
  }
  ~ParallelFastaIngest()
  {
    if (data_) munmap(const_cast<char*>(data_), size_);
    if (fd_ >= 0) close(fd_);
  }
  ParallelFastaIngest(const ParallelFastaIngest&) = delete;
  ParallelFastaIngest& operator=(const ParallelFastaIngest&) = delete;

  bool usable() const { return usable_; }
  void set_window_bytes(uint64_t w) { window_ = std::max<uint64_t>(w, 1); }  // tests shrink it to exercise the seams

  // Streams every record to `sink` in batches of at most cap bytes / max_records records, alternating between
  // buf[0] and buf[1].  Returns the number of records.  Throws std::runtime_error for a record larger than cap.
  uint64_t run(char* const buf[2], size_t cap, size_t max_records, uint64_t first_ordinal, const Sink& sink);

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
  int fd_ = -1;
  const char* data_ = nullptr;
  uint64_t size_ = 0, first_header_ = 0, window_ = 256ull << 20;
  bool usable_ = false;
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
      uint64_t l = lend - q;
      if (l && data_[lend - 1] == '\r') --l;
      r.len += l;
      q = nl ? lend + 1 : size_;
    }
    r.end = q;
    out.push_back(r);
    p = q;
  }
}

inline void ParallelFastaIngest::copy_record(const Rec& r, char* dst) const
{
  uint64_t q = r.seq_begin;
  while (q < r.end) {
    const void* nl = memchr(data_ + q, '\n', r.end - q);
    const uint64_t lend = nl ? uint64_t(static_cast<const char*>(nl) - data_) : r.end;
    uint64_t l = lend - q;
    if (l && data_[lend - 1] == '\r') --l;
    memcpy(dst, data_ + q, l);
    dst += l;
    q = nl ? lend + 1 : r.end;
  }
}

inline uint64_t ParallelFastaIngest::run(char* const buf[2], size_t cap, size_t max_records, uint64_t first_ordinal, const Sink& sink)
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
      const uint64_t wend = std::min(size_, pos + window_);
      const unsigned T = unsigned(std::min<uint64_t>(threads_, std::max<uint64_t>((wend - pos) >> 16, 1)));
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
        if (i == i0) throw std::runtime_error("record larger than the staging buffer in " + path_);
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
