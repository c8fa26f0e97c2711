// links_b200/csrc/host/links_cpp_main.cpp -- `LINKS_CPP`, the C++ host of the B200 hot path.
//
// Drop-in for the reference's LINKS_CPP binary (reference src/main.cpp:21-67 + the stage bodies
// of src/LINKS.hpp it calls): same option letters and defaults (src/InputParser.hpp:68-91,155-255),
// same four timed stages and stdout banners, same output files
//   <b>.bloom  <b>.pairing_issues  <b>.tigpair_checkpoint.tsv
// so bin/LINKS-make (:121-122) and bin/LINKS.pl consume it unchanged.  Every byte of arithmetic on
// the hot path happens behind include/links_gpu.h on the GPU; this file is option parsing, file
// IO and formatting.  There is no CPU fallback: without a CUDA device links_create fails and the
// program exits non-zero.
#include "../../../include/links_gpu.h"
#include "fasta_ingest.hpp"
#include "links_layout.hpp"
#include "seq_stream.hpp"

#include <chrono>
#include <cinttypes>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <getopt.h>
#include <iostream>
#include <sstream>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <vector>

namespace {

const char* const kProgName = "LINKS";
const char* const kVersion = "2.0.1";  // PACKAGE_VERSION of the reference this binary stands in for (configure.ac:1)

struct Options {
  std::string assembly_file, fof_file, base_name, bf_file;
  std::string distances_text, step_sizes_text = "2";
  std::vector<uint32_t> distances{4000};
  std::vector<uint16_t> step_sizes{2};
  uint64_t k = 15, min_size = 500, offset = 0, min_links = 5;
  float insert_stdev = 0.1f, fpr = 0.001f, max_link_ratio = 0.3f;
  bool verbose = false, bf_off = false;
  unsigned thread = 3;
  std::string l_text, a_text, z_text;  // as typed, for the in-process layout stage (LINKS_CPP_LAYOUT=1), which reads them the Perl way
  int device = 0;  // LINKS_GPU_DEVICE environment variable (no new option letter)
  std::vector<std::string> read_files;
  bool ok = true;
};

template <class T> std::vector<T> split_numbers(const std::string& text)
{
  // the reference's own loop shape (src/InputParser.hpp:28-52): `while (ss.good())`, so an empty text and a trailing
  // comma both reach std::stoul("") and end the program with std::invalid_argument there -- and therefore here
  std::vector<T> out;
  std::stringstream ss(text);
  while (ss.good()) {
    std::string item;
    std::getline(ss, item, ',');
    out.push_back(static_cast<T>(std::stoul(item)));
  }
  return out;
}

void usage()
{
  std::cout << "Usage:  " << kProgName << " " << kVersion << "\n"
            << "  -f  sequences to scaffold (Multi-FASTA format, required)\n"
               "  -s  file-of-filenames, full path to long sequence reads (Multi-FASTA/fastq format, required)\n"
               "  -d  distance between k-mer pairs, comma separated (default -d 4000)\n"
               "  -k  k-mer value (default -k 15)\n"
               "  -t  step of sliding window when extracting k-mer pairs, comma separated (default -t 2)\n"
               "  -j  threads: lower bound for the host's FASTA ingest threads (LINKS_HOST_THREADS overrides; no CPU compute workers)\n"
               "  -o  offset position for extracting k-mer pairs (default -o 0)\n"
               "  -e  error (%) allowed on -d distance, e.g. -e 0.1 == distance +/- 10% (default -e 0.1)\n"
               "  -l  minimum number of links to compute scaffold (default -l 5; used by LINKS.pl and the base name)\n"
               "  -a  maximum link ratio between two best contig pairs (default -a 0.3; used by LINKS.pl and the base name)\n"
               "  -z  minimum contig length to consider for scaffolding (default -z 500)\n"
               "  -b  base name for your output files\n"
               "  -r  Bloom filter input file (must derive from the -f file with the same -k)\n"
               "  -p  Bloom filter false positive rate (default -p 0.001)\n"
               "  -x  turn off Bloom filter functionality (-x 1; parsed, and like the reference never consulted)\n"
               "  -v  verbose mode (-v 1)\n";
}

Options parse(int argc, char** argv)
{
  Options o;
  static int help = 0;
  static const struct option longopts[] = {{"help", no_argument, &help, 1}, {nullptr, 0, nullptr, 0}};
  int c, idx = 0;
  while ((c = getopt_long(argc, argv, "f:s:d:k:t:j:o:e:l:a:z:b:r:p:x:v:", longopts, &idx)) != -1) {
    switch (c) {
    case 0: break;
    case 'f': o.assembly_file = optarg; break;
    case 's': o.fof_file = optarg; break;
    case 'd': o.distances_text = optarg; o.distances = split_numbers<uint32_t>(o.distances_text); break;
    case 'k': o.k = strtoul(optarg, nullptr, 10); break;
    case 't': o.step_sizes_text = optarg; o.step_sizes = split_numbers<uint16_t>(o.step_sizes_text); break;
    case 'j': o.thread = unsigned(strtoul(optarg, nullptr, 10)); break;
    case 'o': o.offset = strtoul(optarg, nullptr, 10); break;
    case 'e': o.insert_stdev = strtof(optarg, nullptr); break;
    case 'l': o.min_links = strtoul(optarg, nullptr, 10); o.l_text = optarg; break;
    case 'a': o.max_link_ratio = strtof(optarg, nullptr); o.a_text = optarg; break;
    case 'z': o.min_size = strtoul(optarg, nullptr, 10); o.z_text = optarg; break;
    case 'b': o.base_name = optarg; break;
    case 'r': o.bf_file = optarg; break;
    case 'p': o.fpr = strtof(optarg, nullptr); break;
    case 'x': o.bf_off = strtoul(optarg, nullptr, 10) == 1; break;
    case 'v': o.verbose = strtoul(optarg, nullptr, 10) == 1; break;
    default: exit(EXIT_FAILURE);
    }
  }
  if (help) { usage(); exit(0); }
  // -t is padded with its last value up to the number of distances (InputParser.hpp:224-229)
  while (o.step_sizes.size() < o.distances.size()) o.step_sizes.push_back(o.step_sizes.back());
  {
    std::ifstream fof(o.fof_file);
    std::string path;
    while (fof >> path) o.read_files.push_back(path);
    if (o.read_files.empty()) {
      std::cerr << "\n File of files for reads cannot be empty (-s)\n";
      o.ok = false;
    }
  }
  if (o.distances.size() < o.step_sizes.size()) {
    std::cerr << "\n Number of provided distances can't be lower than number of step sizes provided.\n";
    o.ok = false;
  }
  if (o.base_name.empty()) {
    // default output prefix, same recipe as the reference (InputParser.hpp:241-250)
    o.base_name = o.assembly_file + "scaff_s-" + o.fof_file + "_d" + o.distances_text + "_k" + std::to_string(o.k) + "_e" +
                  std::to_string(o.insert_stdev) + "_l" + std::to_string(o.min_links) + "_a" + std::to_string(o.max_link_ratio) +
                  "_z" + std::to_string(o.min_size) + "_t" + o.step_sizes_text + "_o" + std::to_string(o.offset) + "_r-" +
                  (o.bf_off ? "-" : o.bf_file) + "_p" + std::to_string(o.fpr) + "_x" + std::to_string(o.bf_off);
  }
  if (o.assembly_file.empty() || o.fof_file.empty()) {
    usage();
    o.ok = false;
  }
  if (const char* d = getenv("LINKS_GPU_DEVICE")) o.device = atoi(d);
  return o;
}

void print_opts(const Options& o)
{
  std::cout << "\nLINKS running options:\n"
            << "  -f " << o.assembly_file << "\n  -s " << o.fof_file << "\n  -d " << o.distances_text << "\n  -k " << o.k
            << "\n  -j " << o.thread << "\n  -t " << o.step_sizes_text << "\n  -o " << o.offset << "\n  -e " << o.insert_stdev
            << "\n  -l " << o.min_links << "\n  -a " << o.max_link_ratio << "\n  -z " << o.min_size << "\n  -b " << o.base_name
            << "\n  -r " << o.bf_file << "\n  -p " << o.fpr << "\n  -x " << o.bf_off << "\n  -v " << o.verbose << "\n";
}

std::string format_duration(std::chrono::milliseconds ms)
{
  using namespace std::chrono;
  const auto h = duration_cast<hours>(ms); ms -= h;
  const auto m = duration_cast<minutes>(ms); ms -= m;
  const auto s = duration_cast<seconds>(ms); ms -= s;
  std::ostringstream ss;
  ss << h.count() << "h " << m.count() << "m " << s.count() << "s " << ms.count() << "ms";
  return ss.str();
}

[[noreturn]] void die(links_ctx* ctx, const char* what, int rc)
{
  std::cerr << "LINKS_CPP: " << what << " failed (" << rc << "): " << links_last_error(ctx) << "\n";
  if (ctx) links_destroy(ctx);
  exit(1);
}

#define LK(ctx, call)                         \
  do {                                        \
    const int rc__ = (call);                  \
    if (rc__ != LINKS_OK) die(ctx, #call, rc__); \
  } while (0)

}  // namespace

namespace {

// BTLKmerBloomFilter_v6 file: TOML header, "[HeaderEnd]", 50 newlines with a "<binary data>"
// marker on the second, then the raw bit array (reference btllib bloom_filter.hpp:697-740,766-783).
struct BloomFile { uint32_t k = 0, hash_num = 0; uint64_t bytes = 0; std::vector<uint8_t> bits; };

BloomFile read_bloom_file(const std::string& path)
{
  std::ifstream f(path, std::ios::binary);
  if (!f) { std::cout << "\nInvalid file: " << path << " -- fatal\n"; exit(1); }
  BloomFile bf;
  std::string line;
  bool magic = false, end = false;
  while (std::getline(f, line)) {
    if (line == "[BTLKmerBloomFilter_v6]") { magic = true; continue; }
    if (line == "[HeaderEnd]") { end = true; break; }
    const size_t eq = line.find('=');
    if (eq == std::string::npos) continue;
    std::string key = line.substr(0, eq), val = line.substr(eq + 1);
    auto trim = [](std::string& s) {
      const size_t a = s.find_first_not_of(" \t\""), b = s.find_last_not_of(" \t\"\r");
      s = a == std::string::npos ? "" : s.substr(a, b - a + 1);
    };
    trim(key); trim(val);
    if (key == "k") bf.k = uint32_t(std::stoul(val));
    else if (key == "hash_num") bf.hash_num = uint32_t(std::stoul(val));
    else if (key == "bytes") bf.bytes = std::stoull(val);
  }
  if (!magic || !end || !bf.bytes || !bf.hash_num) {
    std::cerr << "LINKS_CPP: " << path << ": magic string does not match (likely version mismatch)\n";
    exit(1);
  }
  for (int i = 0; i < 50; ++i) std::getline(f, line);  // PLACEHOLDER_NEWLINES
  bf.bits.resize(bf.bytes);
  f.read(reinterpret_cast<char*>(bf.bits.data()), std::streamsize(bf.bytes));
  if (uint64_t(f.gcount()) != bf.bytes) { std::cerr << "LINKS_CPP: " << path << ": truncated Bloom filter file\n"; exit(1); }
  return bf;
}

uint64_t file_size_bytes(const std::string& path)
{
  struct stat st;
  if (stat(path.c_str(), &st) != 0) { std::cerr << "LINKS_CPP: cannot stat " << path << "\n"; exit(1); }
  return uint64_t(st.st_size);
}

struct Stage {
  std::chrono::high_resolution_clock::time_point t0;
  explicit Stage(const char* banner) : t0(std::chrono::high_resolution_clock::now()) { std::cout << banner << "\n"; }
  void done(const char* msg)
  {
    const auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - t0);
    std::cout << msg << format_duration(ms) << "\n\n";
  }
};

// LINKS_CPP_TIMING=1: milliseconds since main() at the host's milestones, on stderr (where the wall time of a short job goes)
void milestone(const char* what)
{
  static const bool on = getenv("LINKS_CPP_TIMING") != nullptr;
  static const auto t0 = std::chrono::steady_clock::now();
  if (on) std::cerr << "[LINKS_CPP] " << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() << " ms  " << what << "\n";
}

}  // namespace

int main(int argc, char** argv)
{
  milestone("main");
  Options o = parse(argc, argv);
  if (!o.ok) return -1;
  print_opts(o);
  std::cout << std::endl;

  links_params prm{};
  prm.k = uint32_t(o.k);
  prm.n_dist = uint32_t(o.distances.size());
  prm.dist = o.distances.data();
  prm.step = o.step_sizes.data();
  prm.offset = o.offset;
  prm.insert_stdev = o.insert_stdev;
  prm.device = o.device;
  links_ctx* ctx = nullptr;
  {
    const int rc = links_create(&ctx, &prm);
    if (rc != LINKS_OK) die(nullptr, "links_create", rc);
  }
  milestone("links_create done (CUDA context, streams, tables, kernels loaded)");

  // two pinned staging buffers shared by the contig and read streams: while one is on its way to the device
  // the ingest threads fill the other.  No larger than the largest input needs (pinning memory is not free).
  size_t stage_mb = 64;  // x2; pinning costs ~0.35 ms per MB, and 64 MB batches already keep the device busy (profiles/r01zb_e2e_binary_c2.json)
  size_t stage_cap = stage_mb << 20;
  if (const char* e = getenv("LINKS_STAGE_MB")) stage_cap = std::max<size_t>(size_t(strtod(e, nullptr) * double(1 << 20)), 4096);  // fractions allowed
  {
    uint64_t need = 0;
    bool known = true;
    std::vector<std::string> inputs = o.read_files;
    inputs.push_back(o.assembly_file);
    for (const std::string& p : inputs) {
      struct stat st;
      if (stat(p.c_str(), &st) != 0 || !S_ISREG(st.st_mode)) { known = false; break; }
      for (const char* suf : {".gz", ".bz2", ".xz", ".zst", ".zip", ".lrz", ".7z"})
        if (links_host::ends_with(p, suf)) known = false;
      need = std::max<uint64_t>(need, uint64_t(st.st_size));
    }
    if (known) stage_cap = std::min<size_t>(stage_cap, size_t(((need + 1) + (1u << 20) - 1) >> 20 << 20));
  }
  unsigned host_threads = std::max(o.thread, std::min(std::thread::hardware_concurrency(), 16u));
  if (const char* e = getenv("LINKS_HOST_THREADS")) host_threads = unsigned(std::max(atoi(e), 1));
  char* stage[2] = {static_cast<char*>(links_host_alloc(stage_cap)), static_cast<char*>(links_host_alloc(stage_cap))};
  if (!stage[0] || !stage[1]) { std::cerr << "LINKS_CPP: cannot allocate 2 x " << (stage_cap >> 20) << " MB of pinned host memory\n"; return 1; }
  milestone("pinned staging buffers allocated");
  links_host::SeqBatch batch;
  batch.cap = stage_cap;
  batch.seq = stage[0];
  // One record longer than a staging buffer (a chromosome-scale contig): both buffers are replaced by larger ones.  Both
  // readers call this with no push in flight; `keep` bytes at the start of stage[0] (the serial reader's open record) survive.
  auto grow_stage = [&](size_t need, size_t keep) -> bool {
    const size_t ncap = ((std::max(need + need / 4, stage_cap * 2) >> 20) + 1) << 20;
    char* a = static_cast<char*>(links_host_alloc(ncap));
    char* b = a ? static_cast<char*>(links_host_alloc(ncap)) : nullptr;
    if (!a || !b) { if (a) links_host_free(a); return false; }
    if (keep) memcpy(a, stage[0], keep);
    links_host_free(stage[0]);
    links_host_free(stage[1]);
    stage[0] = a; stage[1] = b; stage_cap = ncap;
    batch.seq = a; batch.cap = ncap;
    return true;
  };
  batch.grow = [&](links_host::SeqBatch&, size_t need, size_t keep) { return grow_stage(need, keep); };
  // One input file -> batches for push(seq, off, n, first_ordinal); returns the records read.  Plain FASTA goes
  // through the multi-threaded reader; FASTQ, compressed inputs and pipes through the serial
  // one, exactly as before.  A batch never spans two files.
  auto ingest = [&](const std::string& path, size_t max_records, uint64_t first_ordinal,
                    const std::function<void(const char*, const uint64_t*, size_t, uint64_t)>& push) -> uint64_t {
    if (host_threads > 1) {
      links_host::ParallelFastaIngest par(path, host_threads);
      if (par.usable())
        return par.run(stage, stage_cap, max_records, first_ordinal, push,
                       [&](size_t need, char**, size_t&) { return grow_stage(need, 0); });  // (stage / stage_cap are the very objects run() holds)
    }
    batch.clear(first_ordinal);
    return links_host::stream_file(path, batch, max_records, true,
                                   [&](links_host::SeqBatch& b) { push(b.seq, b.off.data(), b.n(), b.first_ordinal); });
  };

  // ---- stage 1: init_bloom_filter (reference src/LINKS.hpp:284-352)
  uint64_t bf_bytes = 0;
  uint32_t hash_num = 0;
  uint64_t n_contigs = 0;
  {
    Stage st("STAGE: Bloom filter construction/read");
    if (!o.bf_file.empty()) {
      std::cout << "A Bloom filter was supplied (" << o.bf_file << ") and will be used instead of building a new one\n";
      const BloomFile bf = read_bloom_file(o.bf_file);
      bf_bytes = bf.bytes; hash_num = bf.hash_num;
      LK(ctx, links_bf_load(ctx, bf.bits.data(), bf.bytes, bf.hash_num));
    } else {
      const uint64_t n = file_size_bytes(o.assembly_file);  // file BYTES, as the reference sizes it (:349)
      uint64_t m = 0;
      LK(ctx, links_bf_size(n, o.fpr, &m, &bf_bytes, &hash_num));
      std::cout << "- Number of BF Elements: " << n << "\n- Input file path: " << o.base_name << "\n- Input file: " << o.assembly_file
                << "\n- kmersize: " << o.k << "\n- m: " << m << "\n- fpr: " << o.fpr << "\n- hashFct: " << hash_num << "\n";
      LK(ctx, links_bf_alloc(ctx, bf_bytes, hash_num));
    }
    // every record of the assembly, no -z filter (:318-324); the packed contigs stay on the
    // device for the k-mer table of stage 3, so the file is read once
    try {
      n_contigs = ingest(o.assembly_file, size_t(1) << 22, 0, [&](const char* seq, const uint64_t* off, size_t n, uint64_t first) {
        LK(ctx, links_contigs_push(ctx, seq, off, uint32_t(n), uint32_t(first + 1)));
      });
    } catch (const std::exception& e) { std::cerr << "LINKS_CPP: " << e.what() << "\n"; return 1; }
    if (o.bf_file.empty()) {
      std::cout << "\n\nWriting Bloom filter to disk (" << o.base_name << ".bloom)\n";
      std::vector<uint8_t> bits(bf_bytes);
      LK(ctx, links_bf_download(ctx, bits.data()));
      char hdr[512];
      const size_t hl = links_bf_header(hdr, sizeof hdr, uint32_t(o.k), hash_num, bf_bytes);
      std::ofstream out(o.base_name + ".bloom", std::ios::binary);
      out.write(hdr, std::streamsize(hl));
      out.write(reinterpret_cast<const char*>(bits.data()), std::streamsize(bf_bytes));
      if (!out) { std::cerr << "LINKS_CPP: cannot write " << o.base_name << ".bloom\n"; return 1; }
    }
    st.done("Bloom filter generated or loaded in: ");
    milestone("stage 1 done");
  }

  // ---- stage 2: start_read_fasta (:470-486): every file of the .fof in order; read ordinals run
  // across files (they define which distance wins when several -d see the same k-mer pair)
  {
    Stage st("STAGE: Reading sequence reads");
    uint64_t n_reads = 0;
    try {
      for (const std::string& path : o.read_files) {
        n_reads += ingest(path, size_t(1) << 24, n_reads, [&](const char* seq, const uint64_t* off, size_t n, uint64_t first) {
          LK(ctx, links_reads_push(ctx, seq, off, uint32_t(n), first));
          if (o.verbose) std::cout << "Processed " << first + n << " reads\n";
        });
      }
    } catch (const std::exception& e) { std::cerr << "LINKS_CPP: " << e.what() << "\n"; return 1; }
    LK(ctx, links_sync(ctx));
    std::cout << "Reads processed: " << n_reads << "\n";
    st.done("Read fasta stage completed in: ");
    milestone("stage 2 done");
  }

  // ---- stage 3: start_read_contig (:487-513).  The contigs were packed and kept resident by
  // stage 1; the k-mer table itself is the first kernel of links_finalize.
  {
    Stage st("STAGE: Reading contigs");
    std::cout << "Contigs resident on device: " << n_contigs << "\n";
    st.done("Read contig stage completed in: ");
  }

  // ---- stage 4: pair_contigs (:780-1041)
  {
    Stage st("STAGE: Pairing contigs");
    links_row* rows = nullptr; links_issue* issues = nullptr;
    uint64_t n_rows = 0, n_issues = 0;
    links_stats s{};
    LK(ctx, links_finalize(ctx, &rows, &n_rows, &issues, &n_issues, &s));
    {
      std::ofstream f(o.base_name + ".pairing_issues", std::ios::binary);
      char line[1024];
      for (uint64_t i = 0; i < n_issues; ++i) { const int n = links_format_issue(&issues[i], line, sizeof line); f.write(line, n); }
    }
    const uint64_t satisfied = s.ct_ok_pairs + s.ct_ok_contig, unsatisfied = s.ct_problem_pairs + s.ct_iz_issues + s.ct_illogical;
    std::cout << "\n===========PAIRED K-MER STATS===========\n"
              << "Candidate k-mer pairs visited: " << s.n_candidates << "\n"
              << "Pairs with both k-mers in the Bloom filter: " << s.n_survivors << " (" << s.n_unique_pairs << " distinct)\n"
              << "Pairs with an ambiguous (multi-copy) k-mer: " << s.ct_gate_fail << "\n"
              << "\tSatisfied in distance/logic within contigs (i.e. -> <-, distance on target): " << s.ct_ok_contig << "\n"
              << "\tUnsatisfied in distance within contigs (i.e. distance out-of-bounds): " << s.ct_iz_issues << "\n"
              << "\tUnsatisfied pairing logic within contigs (i.e. illogical pairing ->->, <-<- or <-->): " << s.ct_illogical << "\n"
              << "\t---\n"
              << "\tSatisfied in distance/logic within a given contig pair (pre-scaffold): " << s.ct_ok_pairs << "\n"
              << "\tUnsatisfied in distance within a given contig pair (i.e. calculated distances out-of-bounds): " << s.ct_problem_pairs << "\n"
              << "\t---\n"
              << "Total satisfied: " << satisfied << "\tunsatisfied: " << unsatisfied
              << "\n============================================\n\n";
    {
      std::ofstream f(o.base_name + ".tigpair_checkpoint.tsv", std::ios::binary);
      char line[256];
      for (uint64_t i = 0; i < n_rows; ++i) { const int n = links_format_row(&rows[i], line, sizeof line); f.write(line, n); }
      if (!f) { std::cerr << "LINKS_CPP: cannot write " << o.base_name << ".tigpair_checkpoint.tsv\n"; return 1; }
    }
    st.done("Pair contig stage completed in: ");
    if (getenv("LINKS_CPP_LAYOUT")) {
      // the layout stage (bin/LINKS.pl, here links_layout.hpp) on the rows still in memory -- same
      // files as `LINKS.pl -f <f> -b <b> -l -a -z` run afterwards would write, without reading the tsv back
      std::vector<links_layout::RowIn> in(n_rows);
      for (uint64_t i = 0; i < n_rows; ++i)
        in[i] = links_layout::RowIn{rows[i].bin, rows[i].tig_a, rows[i].tig_b, rows[i].rev_a != 0, rows[i].rev_b != 0, rows[i].links, rows[i].gaps};
      links_layout::Options lo;
      lo.assembly_file = o.assembly_file;
      lo.base_name = o.base_name;
      if (links_layout::perl_true(o.l_text)) lo.min_links = strtoll(o.l_text.c_str(), nullptr, 10);
      if (links_layout::perl_true(o.z_text)) lo.min_size = strtoll(o.z_text.c_str(), nullptr, 10);
      if (links_layout::perl_true(o.a_text)) lo.max_link_ratio = strtod(o.a_text.c_str(), nullptr);
      Stage ls("STAGE: Scaffold layout (in process)");
      const int lrc = links_layout::run(lo, stdout, in.data(), in.size());
      ls.done("Layout stage completed in: ");
      if (lrc) { links_free_result(rows, issues); return lrc; }
    }
    links_free_result(rows, issues);
    milestone("stage 4 done");
  }

  links_host_free(stage[0]);
  links_host_free(stage[1]);
  milestone("staging freed");
  links_destroy(ctx);
  milestone("links_destroy done");
  return 0;
}
