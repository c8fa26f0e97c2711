// tests/host/ingest_check.cpp -- the multi-threaded FASTA ingest (links_b200/csrc/host/fasta_ingest.hpp) against the
// serial reader (seq_stream.hpp) on the same file: same records, same order, same ordinals.  CPU only.
//   ingest_check <file> <threads> <cap_bytes> <window_bytes> <max_records> [min_segment_bytes] [grow]
// with `grow` both readers may enlarge their buffers for a record that does not fit (as LINKS_CPP lets them)
// prints "OK <records> <bases>" / "FALLBACK" (file is not plain FASTA) / "ERROR <what>" / "MISMATCH ..."
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "fasta_ingest.hpp"
#include "seq_stream.hpp"

int main(int argc, char** argv)
{
  if (argc < 6) return 2;
  const std::string path = argv[1];
  const unsigned threads = unsigned(atoi(argv[2]));
  const bool allow_grow = argc > 7 && std::string(argv[7]) == "grow";
  const size_t cap = size_t(atoll(argv[3])), window = size_t(atoll(argv[4])), max_records = size_t(atoll(argv[5]));
  std::vector<std::string> serial, par;
  std::vector<uint64_t> serial_ord, par_ord;
  std::string serial_err, par_err;
  {
    std::vector<char> buf(cap);
    links_host::SeqBatch b;
    b.seq = buf.data();
    b.cap = cap;
    b.clear(7);
    if (allow_grow)
      b.grow = [&](links_host::SeqBatch& x, size_t need, size_t keep) {
        std::vector<char> bigger(need * 2);
        memcpy(bigger.data(), x.seq, keep);
        buf.swap(bigger);
        x.seq = buf.data();
        x.cap = buf.size();
        return true;
      };
    try {
      links_host::stream_file(path, b, max_records, true, [&](links_host::SeqBatch& x) {
        for (size_t i = 0; i < x.n(); ++i) {
          serial.emplace_back(x.seq + x.off[i], x.seq + x.off[i + 1]);
          serial_ord.push_back(x.first_ordinal + i);
        }
      });
    } catch (const std::exception& e) { serial_err = e.what(); }
  }
  links_host::ParallelFastaIngest ing(path, threads, argc > 6 ? uint64_t(atoll(argv[6])) : (1u << 16));
  if (!ing.usable()) {  // optional detail for tests that care what the serial reader then did
    if (getenv("INGEST_CHECK_VERBOSE")) {
      uint64_t bases = 0;
      for (auto& r : serial) bases += r.size();
      if (serial_err.empty()) printf("FALLBACK %zu %llu\n", serial.size(), (unsigned long long)bases);
      else printf("FALLBACK ERROR\n");
    } else printf("FALLBACK\n");
    return 0;
  }
  ing.set_window_bytes(window);
  std::vector<char> b0(cap), b1(cap);
  char* bufs[2] = {b0.data(), b1.data()};
  uint64_t n = 0;
  size_t pcap = cap;
  links_host::ParallelFastaIngest::Grow grow;
  if (allow_grow)
    grow = [&](size_t need, char* bb[2], size_t& c) {
      b0.assign(need * 2, 0); b1.assign(need * 2, 0);
      bb[0] = b0.data(); bb[1] = b1.data(); c = need * 2;
      return true;
    };
  try {
    n = ing.run(bufs, pcap, max_records, 7, [&](const char* seq, const uint64_t* off, size_t cnt, uint64_t first) {
      for (size_t i = 0; i < cnt; ++i) {
        par.emplace_back(seq + off[i], seq + off[i + 1]);
        par_ord.push_back(first + i);
      }
    }, grow);
  } catch (const std::exception& e) { par_err = e.what(); }
  if (!serial_err.empty() || !par_err.empty()) {
    if (!serial_err.empty() && !par_err.empty()) { printf("ERROR both\n"); return 0; }
    printf("MISMATCH errors: serial='%s' parallel='%s'\n", serial_err.c_str(), par_err.c_str());
    return 1;
  }
  if (n != par.size() || par.size() != serial.size()) { printf("MISMATCH counts %zu %zu %llu\n", serial.size(), par.size(), (unsigned long long)n); return 1; }
  uint64_t bases = 0;
  for (size_t i = 0; i < par.size(); ++i) {
    if (par[i] != serial[i] || par_ord[i] != serial_ord[i]) { printf("MISMATCH record %zu (len %zu vs %zu)\n", i, serial[i].size(), par[i].size()); return 1; }
    bases += par[i].size();
  }
  printf("OK %zu %llu\n", par.size(), (unsigned long long)bases);
  return 0;
}
