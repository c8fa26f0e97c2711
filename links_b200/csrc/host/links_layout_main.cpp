// links_b200/csrc/host/links_layout_main.cpp -- command line of the layout stage: a drop-in for the reference's
// bin/LINKS.pl as LINKS-make calls it (bin/LINKS-make:124: `LINKS.pl -f draft -s empty.fof -b b -l l -a a -e e -z z -p p -v v`).
// Installed as links_b200/bin/LINKS_layout and, for the unmodified LINKS-make, under the name LINKS.pl.
#include "links_layout.hpp"

#include <unistd.h>

int main(int argc, char** argv)
{
  links_layout::Options o;
  std::string opt_l, opt_a, opt_z, opt_v;
  int ch;
  // same letters as getopts('f:s:d:k:e:l:a:v:b:t:p:o:r:x:m:z:') (bin/LINKS.pl:27); the ones the layout ignores are accepted
  while ((ch = getopt(argc, argv, "f:s:d:k:e:l:a:v:b:t:p:o:r:x:m:z:")) != -1) {
    switch (ch) {
    case 'f': o.assembly_file = optarg; break;
    case 'b': o.base_name = optarg; break;
    case 'l': opt_l = optarg; break;
    case 'a': opt_a = optarg; break;
    case 'z': opt_z = optarg; break;
    case 'v': opt_v = optarg; break;
    default: break;
    }
  }
  if (o.assembly_file.empty() || o.base_name.empty()) {
    fprintf(stderr, "Usage: %s %s\n-f  sequences to scaffold (Multi-FASTA format, required)\n-b  base name of the LINKS_CPP run (<b>.tigpair_checkpoint.tsv is read, required)\n"
                    "-l  minimum number of links (k-mer pairs) to compute scaffold (default -l 5, optional)\n"
                    "-a  maximum link ratio between two best contig pairs (default -a 0.3, optional)\n"
                    "-z  minimum contig length to consider for scaffolding (default -z 500, optional)\n"
                    "-v  verbose (accepted)\n\nError: Missing mandatory options -f and -b.\n\n", argv[0], o.version.c_str());
    return 2;
  }
  // `$min_links = $opt_l if($opt_l)`: a value Perl finds false ("", "0") leaves the default (bin/LINKS.pl:52-55)
  if (links_layout::perl_true(opt_l)) o.min_links = strtoll(opt_l.c_str(), nullptr, 10);
  if (links_layout::perl_true(opt_z)) o.min_size = strtoll(opt_z.c_str(), nullptr, 10);
  if (links_layout::perl_true(opt_a)) o.max_link_ratio = strtod(opt_a.c_str(), nullptr);
  if (links_layout::perl_true(opt_v)) o.verbose = atoi(opt_v.c_str());
  if (access(o.assembly_file.c_str(), R_OK) != 0) { fprintf(stderr, "Invalid file: %s -- fatal\n", o.assembly_file.c_str()); return 1; }
  printf("\nRunning: %s %s\n-f %s\n-l %lld\n-a %g\n-z %lld\n-b %s\n\n", argv[0], o.version.c_str(), o.assembly_file.c_str(), o.min_links,
         o.max_link_ratio, o.min_size, o.base_name.c_str());
  const int rc = links_layout::run(o);
  printf("%s %s terminated successfully\n", argv[0], o.version.c_str());
  return rc;
}
