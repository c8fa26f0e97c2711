// Test driver for links_b200/csrc/host/seq_stream.hpp: prints "<ordinal>\t<length>\t<fnv1a>" per
// record of the files given, streaming through a deliberately small staging buffer so the
// batch-rollover path is exercised.  usage: seq_stream_dump <cap_bytes> <max_records> file...
#include "../../links_b200/csrc/host/seq_stream.hpp"

#include <cinttypes>
#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv)
{
  if (argc < 4) return 2;
  links_host::SeqBatch b;
  b.cap = size_t(atoll(argv[1]));
  std::vector<char> buf(b.cap);
  b.seq = buf.data();
  b.clear(0);
  auto dump = [](links_host::SeqBatch& x) {
    for (size_t i = 0; i < x.n(); ++i) {
      uint64_t h = 1469598103934665603ull;
      for (uint64_t j = x.off[i]; j < x.off[i + 1]; ++j) { h ^= (unsigned char)x.seq[j]; h *= 1099511628211ull; }
      printf("%" PRIu64 "\t%" PRIu64 "\t%016" PRIx64 "\n", x.first_ordinal + i, x.off[i + 1] - x.off[i], h);
    }
  };
  try {
    for (int a = 3; a < argc; ++a) links_host::stream_file(argv[a], b, size_t(atoll(argv[2])), false, dump);
    if (b.n()) dump(b);
  } catch (const std::exception& e) { fprintf(stderr, "%s\n", e.what()); return 1; }
  return 0;
}
