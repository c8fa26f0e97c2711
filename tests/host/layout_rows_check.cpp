// tests/host/layout_rows_check.cpp -- the layout stage fed from rows in memory (links_layout::run with RowIn, what
// LINKS_CPP does under LINKS_CPP_LAYOUT=1) against the same stage reading the tsv those rows print as.  CPU only.
//   layout_rows_check <contigs.fa> <rows.tsv> <base_file> <base_mem> <l> <a> <z>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <vector>

#include "links_layout.hpp"

int main(int argc, char** argv)
{
  if (argc < 8) return 2;
  links_layout::Options o;
  o.assembly_file = argv[1];
  o.min_links = atoll(argv[5]);
  o.max_link_ratio = strtod(argv[6], nullptr);
  o.min_size = atoll(argv[7]);
  std::vector<links_layout::RowIn> rows;
  {
    std::ifstream in(argv[2]);
    std::ofstream copy(std::string(argv[3]) + ".tigpair_checkpoint.tsv");
    std::string line;
    while (std::getline(in, line)) {
      copy << line << "\n";
      std::istringstream ss(line);
      std::string d, a, b;
      unsigned long long links; long long gaps;
      ss >> d >> a >> b >> links >> gaps;
      rows.push_back(links_layout::RowIn{atoi(d.c_str()), uint32_t(strtoul(a.c_str() + 1, nullptr, 10)), uint32_t(strtoul(b.c_str() + 1, nullptr, 10)),
                                         a[0] == 'r', b[0] == 'r', links, gaps});
    }
  }
  FILE* devnull = fopen("/dev/null", "w");
  o.base_name = argv[3];
  if (links_layout::run(o, devnull)) return 1;
  o.base_name = argv[4];
  if (links_layout::run(o, devnull, rows.data(), rows.size())) return 1;
  return 0;
}
