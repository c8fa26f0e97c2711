// links_b200/csrc/host/links_layout.hpp -- the scaffold layout stage of LINKS (CPU, C++).
//
// Restates the second half of the reference pipeline, bin/LINKS.pl (what `LINKS-make` runs after LINKS_CPP:
// `LINKS.pl -f draft -s empty.fof -b base -l -a -e -z -p -v`), function by function:
//   read_contigs_memory   readContigsMemory   bin/LINKS.pl:1238-1303
//   read_contig_pairs     readContigPairs     bin/LINKS.pl:1146-1235
//   build_scaffolds       buildScaffolds      bin/LINKS.pl:662-700
//   compute_layout        computeLayout       bin/LINKS.pl:703-809
//   build_graph           buildGraph          bin/LINKS.pl:620-650
//   build_scaffold_fasta  buildScaffoldFasta  bin/LINKS.pl:1306-1382 (+ reverseComplement :653-658)
// Same files out: <b>.scaffolds, <b>.scaffolds.fa, <b>.gv, <b>.assembly_correspondence.tsv, <b>.log.
// SURVEY.md 8(f) rows 2-3.  Not on the GPU path: the layout is a greedy walk over a few thousand rows.
//
// Where the reference is not deterministic, this is: Perl iterates hashes in a per-process random order, so
// LINKS.pl breaks TIES (contigs of equal length as seeds; two neighbours with equal link counts) differently
// from run to run.  Here ties go to the smaller contig id / the lexicographically smaller neighbour name.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace links_layout {

struct Options {
  std::string assembly_file, base_name;
  long long min_links = 5;      // -l
  double max_link_ratio = 0.3;  // -a
  long long min_size = 500;     // -z
  int verbose = 0;
  std::string version = "[v1.8.6]";  // bin/LINKS.pl:32 (the Perl stage kept its own version string)
};

struct Contigs {  // ids are 1-based ordinals (readContigsMemory numbers the records it keeps)
  std::vector<std::string> name, seq;  // [0] unused
  uint64_t length(long long id) const { return id >= 1 && size_t(id) < seq.size() ? seq[size_t(id)].size() : 0; }
  long long count() const { return (long long)seq.size() - 1; }
};

struct LinkGap { long long links = 0, gaps = 0; };
// pair{tig}{'layer'}{distance}{neighbour} ; distance keys keep their text (they sort numerically, :735)
struct Layer { std::string key; double num; std::map<std::string, LinkGap> list; };
struct Node { std::vector<Layer> layers; };  // sorted by num on use
using PairMap = std::unordered_map<std::string, Node>;

struct Simple {  // simplepair{tig1}{tig2}
  bool has_links = false; long long links = 0, sumgaps = 0, max = 0; std::string type;
  bool linked = false; long long mean = 0;
};
using SimpleMap = std::map<long long, std::map<long long, Simple>>;

inline FILE* open_maybe_gz(const std::string& path, bool& piped)
{
  auto ends = [&](const char* suf) {
    const size_t n = strlen(suf);
    if (path.size() < n) return false;
    for (size_t i = 0; i < n; ++i) if (tolower(path[path.size() - n + i]) != suf[i]) return false;
    return true;
  };
  std::string quoted = "'";
  for (char ch : path) { if (ch == '\'') quoted += "'\\''"; else quoted += ch; }
  quoted += "'";
  piped = true;
  if (ends("zip")) return popen(("unzip -p " + quoted).c_str(), "r");                       // :1251
  if (ends("gz") || ends("gzip")) return popen(("gunzip -c " + quoted).c_str(), "r");       // :1253
  piped = false;
  return fopen(path.c_str(), "r");
}

inline bool read_line(FILE* f, std::string& line)  // chomp semantics: strips one trailing '\n'
{
  line.clear();
  char buf[1 << 16];
  bool any = false;
  while (fgets(buf, sizeof buf, f)) {
    any = true;
    const size_t n = strlen(buf);
    if (n && buf[n - 1] == '\n') { line.append(buf, n - 1); return true; }
    line.append(buf, n);
  }
  return any;
}

// bin/LINKS.pl:1280-1300: a record is kept when its header differs from the previous one, the previous record
// had sequence and a header; the last record is always kept.  Sequence is upper-cased.
inline Contigs read_contigs_memory(const std::string& file)
{
  bool piped = false;
  FILE* in = open_maybe_gz(file, piped);
  if (!in) throw std::runtime_error("Error reading " + file + " -- fatal");
  Contigs c;
  c.name.emplace_back(); c.seq.emplace_back();
  std::string line, prevhead, seq;
  while (read_line(in, line)) {
    if (!line.empty() && line[0] == '>') {
      const std::string head = line.substr(1);
      if (head != prevhead && !seq.empty() && !prevhead.empty()) { c.name.push_back(prevhead); c.seq.push_back(seq); }
      seq.clear();
      prevhead = head;
    } else {
      for (char& ch : line) ch = char(toupper((unsigned char)ch));
      seq += line;
    }
  }
  c.name.push_back(prevhead); c.seq.push_back(seq);
  if (piped) pclose(in); else fclose(in);
  return c;
}

inline std::vector<std::string> split_tab(const std::string& s)
{
  std::vector<std::string> out;
  size_t a = 0;
  for (;;) {
    const size_t b = s.find('\t', a);
    if (b == std::string::npos) { out.push_back(s.substr(a)); break; }
    out.push_back(s.substr(a, b - a));
    a = b + 1;
  }
  while (!out.empty() && out.back().empty()) out.pop_back();  // Perl's split drops trailing empty fields
  return out;
}

inline bool perl_true(const std::string& s) { return !s.empty() && s != "0"; }
inline long long to_ll(const std::string& s) { return s.empty() ? 0 : strtoll(s.c_str(), nullptr, 10); }

// "f12" -> ('f', 12); false when the text is not [rf]\d+
inline bool parse_tig(const std::string& s, char& dir, long long& id)
{
  if (s.size() < 2 || (s[0] != 'f' && s[0] != 'r')) return false;
  for (size_t i = 1; i < s.size(); ++i) if (s[i] < '0' || s[i] > '9') return false;
  dir = s[0]; id = strtoll(s.c_str() + 1, nullptr, 10);
  return true;
}

inline Layer& layer_of(Node& n, const std::string& key)
{
  for (auto& l : n.layers) if (l.key == key) return l;
  n.layers.push_back(Layer{key, strtod(key.c_str(), nullptr), {}});
  return n.layers.back();
}

// One row of <b>.tigpair_checkpoint.tsv as readContigPairs takes it (bin/LINKS.pl:1159-1215): dist, "f12", "r7", links, gaps
inline void add_pair_row(const std::string& dist, const std::string& f1, const std::string& f2, long long links, long long gaps,
                         bool simple_flag, const Contigs& tigs, const Options& o, PairMap& pair, SimpleMap& simple)
{
  // contigs shorter than -z are skipped; an unknown or empty id has length 0 (:1166-1168)
  const std::string uid = f1.empty() ? "" : f1.substr(1), vid = f2.empty() ? "" : f2.substr(1);
  auto len_of = [&](const std::string& id) -> long long {
    if (id.empty()) return 0;
    for (char ch : id) if (ch < '0' || ch > '9') return 0;
    return (long long)tigs.length(strtoll(id.c_str(), nullptr, 10));
  };
  if (len_of(uid) < o.min_size || len_of(vid) < o.min_size) return;
  if (!perl_true(f2)) return;
  LinkGap& lg = layer_of(pair[f1], dist).list[f2];
  lg.links += links; lg.gaps += gaps;
  if (simple_flag) {
    char d1, d2; long long t1, t2;
    if (parse_tig(f1, d1, t1) && parse_tig(f2, d2, t2) && t1 < t2) {
      Simple& s = simple[t1][t2];
      s.has_links = true; s.links += links; s.sumgaps += gaps;
      if (links >= s.max) {  // direction with most support (:1207-1210)
        s.type = std::string(d1 == 'f' ? "1" : "0") + (d2 == 'f' ? "1" : "0");
        s.max = links;
      }
    }
  }
}

inline void read_simple_pairs(const std::string& simplefile, SimpleMap& simple)  // optional <b>.simplepair_checkpoint.tsv (:1218-1231)
{
  FILE* spr = fopen(simplefile.c_str(), "r");
  if (!spr) return;
  std::string line;
  while (read_line(spr, line)) {
    const std::vector<std::string> info = split_tab(line);
    if (info.size() > 2 && perl_true(info[2])) {
      Simple& s = simple[to_ll(info[0])][to_ll(info[1])];
      s.has_links = true; s.links = to_ll(info[2]); s.sumgaps = info.size() > 3 ? to_ll(info[3]) : 0;
      s.type = info.size() > 4 ? info[4] : "";
    }
  }
  fclose(spr);
}

// bin/LINKS.pl:1146-1235
inline void read_contig_pairs(const std::string& pairfile, const std::string& simplefile, const Contigs& tigs, const Options& o,
                              PairMap& pair, SimpleMap& simple)
{
  bool simple_flag = true;
  if (FILE* t = fopen(simplefile.c_str(), "r")) { simple_flag = false; fclose(t); }  // an existing file is not overridden
  FILE* in = fopen(pairfile.c_str(), "r");
  if (!in) throw std::runtime_error("Can't read " + pairfile + " -- fatal.");
  std::string line;
  while (read_line(in, line)) {
    const std::vector<std::string> info = split_tab(line);
    const std::string f1 = info.size() > 1 ? info[1] : "", f2 = info.size() > 2 ? info[2] : "";
    add_pair_row(info[0], f1, f2, info.size() > 3 ? to_ll(info[3]) : 0, info.size() > 4 ? to_ll(info[4]) : 0, simple_flag, tigs, o, pair,
                 simple);
  }
  fclose(in);
  read_simple_pairs(simplefile, simple);
}

// The same from rows still in memory (no round trip through the tsv).  A row is what
// links_finalize returns (include/links_gpu.h: links_row) and what links_format_row prints as one tsv line.
struct RowIn { int bin; uint32_t tig_a, tig_b; bool rev_a, rev_b; unsigned long long links; long long gaps; };
inline void pairs_from_rows(const RowIn* rows, size_t n, const std::string& simplefile, const Contigs& tigs, const Options& o,
                            PairMap& pair, SimpleMap& simple)
{
  bool simple_flag = true;
  if (FILE* t = fopen(simplefile.c_str(), "r")) { simple_flag = false; fclose(t); }
  for (size_t i = 0; i < n; ++i) {
    const RowIn& r = rows[i];
    // contig id 0 prints as an empty id, as links_format_row does
    const std::string a = r.tig_a ? std::to_string(r.tig_a) : std::string(), b = r.tig_b ? std::to_string(r.tig_b) : std::string();
    add_pair_row(std::to_string(r.bin), (r.rev_a ? "r" : "f") + a, (r.rev_b ? "r" : "f") + b, (long long)r.links, r.gaps, simple_flag, tigs,
                 o, pair, simple);
  }
  read_simple_pairs(simplefile, simple);
}

// Perl prints a number with %.15g; `if ($ratio =~ /(\d+\.\d{2})/){$ratio = $1;}` (:757) then keeps the first
// "digits, dot, two digits" it finds in that text (0.3333... -> "0.33", 0.5 stays "0.5", 1.25e-05 -> "1.25")
inline std::string perl_ratio_text(double ratio)
{
  char buf[64];
  snprintf(buf, sizeof buf, "%.15g", ratio);
  const std::string s = buf;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] < '0' || s[i] > '9') continue;
    size_t j = i;
    while (j < s.size() && s[j] >= '0' && s[j] <= '9') ++j;
    // Perl's \d+ backtracks, but the leftmost match always starts at the first digit of a run
    if (j + 2 < s.size() && s[j] == '.' && isdigit((unsigned char)s[j + 1]) && isdigit((unsigned char)s[j + 2])) return s.substr(i, j + 3 - i);
    i = j;
  }
  return s;
}

struct ChainStep {  // one contig of a scaffold, left to right
  char orient; long long tig; uint64_t size;
  bool has_link = false; long long links = 0; std::string ratio; long long mean = 0;  // k, a, m of the join to the RIGHT
};

struct LayoutState {
  PairMap pair;
  SimpleMap simple;
  std::set<long long> seen_it;
};

// bin/LINKS.pl:703-809.  ext 'R': appends to `chain`; 'L': prepends (orientation flipped, :783-786).
inline void compute_layout(char ext, std::vector<ChainStep>& chain, std::string tig, LayoutState& st, const Contigs& tigs,
                           uint64_t& total, long long orig_tig_number, const Options& o)
{
  bool extension = true;
  while (extension) {
    char tdir; long long tnum = 0;
    parse_tig(tig, tdir, tnum);
    auto it = st.pair.find(tig);
    if (it == st.pair.end() || it->second.layers.empty()) break;  // NO MORE MATCH
    std::vector<const Layer*> order;
    for (const Layer& l : it->second.layers) order.push_back(&l);
    std::stable_sort(order.begin(), order.end(), [](const Layer* a, const Layer* b) { return a->num < b->num; });
    bool extended = false;
    for (const Layer* layer : order) {
      // best and second best neighbour by links (:742-753); ties: smaller name first (see the header)
      std::string match1, match2;
      long long link1 = 0, gaps1 = 0, link2 = 0;
      {
        std::vector<std::pair<std::string, LinkGap>> v(layer->list.begin(), layer->list.end());
        std::stable_sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.second.links > b.second.links; });
        if (!v.empty()) { match1 = v[0].first; link1 = v[0].second.links; gaps1 = v[0].second.gaps; }
        if (v.size() > 1) { match2 = v[1].first; link2 = v[1].second.links; }
      }
      double ratio = 0.0;
      if (link1) ratio = double(link2) / double(link1);
      const std::string ratio_text = perl_ratio_text(ratio);
      const double ratio_num = strtod(ratio_text.c_str(), nullptr);
      long long mean = 0;
      if (link1) mean = (long long)(double(gaps1) / double(link1));  // int(): truncation toward zero
      char mdir = 'f'; long long tempnum = 0;
      parse_tig(match1, mdir, tempnum);
      if (st.seen_it.count(tempnum) || link1 < o.min_links || ratio_num > o.max_link_ratio || tempnum == orig_tig_number) {
        extension = false;
        continue;  // next DISTANCE
      }
      const long long o1 = std::min(tnum, tempnum), o2 = std::max(tnum, tempnum);
      Simple& sp = st.simple[o1][o2];
      sp.linked = true; sp.mean = mean;
      const uint64_t tlen = tigs.length(tempnum);
      if (ext == 'R') {
        chain.back().has_link = true; chain.back().links = link1; chain.back().ratio = ratio_text; chain.back().mean = mean;
        chain.push_back(ChainStep{mdir, tempnum, tlen});
      } else {
        ChainStep s{mdir == 'r' ? 'f' : 'r', tempnum, tlen};
        s.has_link = true; s.links = link1; s.ratio = ratio_text; s.mean = mean;
        chain.insert(chain.begin(), s);
      }
      total += tlen;
      tig = match1;
      extension = true;
      st.seen_it.insert(tnum);
      st.seen_it.insert(tempnum);
      extended = true;
      break;  // last DISTANCE: move to the next contig
    }
    if (!extended) break;
  }
}

struct Scaffold { long long number; uint64_t total; std::vector<ChainStep> chain; long long seed; std::string text; };

inline std::string chain_text(const std::vector<ChainStep>& chain, long long seed)
{
  std::string s;
  for (size_t i = 0; i < chain.size(); ++i) {
    const ChainStep& c = chain[i];
    if (i) s += "_";
    s += c.orient; s += std::to_string(c.tig);
    s += (c.tig == seed && c.orient == 'f') ? "Z" : "z";  // the seed is written with a capital Z (:681)
    s += std::to_string(c.size);
    if (c.has_link) s += "k" + std::to_string(c.links) + "a" + c.ratio + "m" + std::to_string(c.mean);
  }
  return s;
}

// bin/LINKS.pl:662-700
inline std::vector<Scaffold> build_scaffolds(LayoutState& st, const Contigs& tigs, const Options& o, const std::string& scaffold_file)
{
  FILE* sc = fopen(scaffold_file.c_str(), "w");
  if (!sc) throw std::runtime_error("Can't write to " + scaffold_file + " -- fatal");
  std::vector<long long> seeds;
  for (long long t = 1; t <= tigs.count(); ++t) seeds.push_back(t);
  std::stable_sort(seeds.begin(), seeds.end(), [&](long long a, long long b) { return tigs.length(a) > tigs.length(b); });
  std::vector<Scaffold> out;
  long long sc_ct = 0;
  for (long long tig : seeds) {
    if (st.seen_it.count(tig)) continue;
    ++sc_ct;
    const std::string ftig = "f" + std::to_string(tig), rtig = "r" + std::to_string(tig);
    uint64_t total = tigs.length(tig);
    std::vector<ChainStep> right{ChainStep{'f', tig, tigs.length(tig)}}, left;
    compute_layout('R', right, ftig, st, tigs, total, tig, o);
    // the left walk starts from the seed's reverse strand; its joins are recorded on the step they precede
    std::vector<ChainStep> walk{ChainStep{'f', tig, tigs.length(tig)}};
    compute_layout('L', walk, rtig, st, tigs, total, tig, o);
    walk.pop_back();  // what is left are the prepended contigs, leftmost first
    left = walk;
    st.pair.erase(ftig); st.pair.erase(rtig);
    st.seen_it.insert(tig);
    Scaffold s;
    s.number = sc_ct; s.total = total; s.seed = tig;
    s.chain = left; s.chain.insert(s.chain.end(), right.begin(), right.end());
    s.text = chain_text(s.chain, tig);
    fprintf(sc, "scaffold%lld,%llu,%s\n", sc_ct, (unsigned long long)total, s.text.c_str());
    out.push_back(std::move(s));
  }
  fclose(sc);
  return out;
}

inline std::string fmt1(double v) { char b[64]; snprintf(b, sizeof b, "%.1f", v); return b; }

// bin/LINKS.pl:620-650 -- the label line carries `date`'s output, so .gv files only compare below that line
inline void build_graph(const SimpleMap& simple, const Contigs& tigs, const Options& o, const std::string& graph_file, const std::string& date)
{
  FILE* gv = fopen(graph_file.c_str(), "w");
  if (!gv) throw std::runtime_error("Can't write to " + graph_file + " -- fatal");
  fprintf(gv, "graph LINKS{\n");
  fprintf(gv, "\tlabel=\"LINKS %s %s %s;\"\n\trankdir=LR;\n\tnode [shape = circle];\n", o.version.c_str(), o.base_name.c_str(), date.c_str());
  for (const auto& a : simple) {
    for (const auto& b : a.second) {
      const Simple& s = b.second;
      if (!(s.links >= o.min_links)) continue;  // only what passes -l is drawn (:630)
      const std::string t1 = std::to_string(a.first), t2 = std::to_string(b.first);
      const std::string s1 = fmt1(double(tigs.length(a.first)) / 1000.0), s2 = fmt1(double(tigs.length(b.first)) / 1000.0);
      if (s.linked) {
        fprintf(gv, "\t%s [style=filled, fillcolor=deepskyblue, color=deepskyblue]\n\t%s [style=filled, fillcolor=deepskyblue, color=deepskyblue]\n",
                t1.c_str(), t2.c_str());
        fprintf(gv, "\t%s -- %s [ label = \"%s_%skb - %s_%skb l=%lld g=%lld type=%s\", penwidth=2.0, color=deepskyblue ]\n", t1.c_str(), t2.c_str(),
                t1.c_str(), s1.c_str(), t2.c_str(), s2.c_str(), s.links, s.mean, s.type.c_str());
      } else {
        const std::string avg = s.links ? fmt1(double(s.sumgaps) / double(s.links)) : "";
        fprintf(gv, "\t%s -- %s [ label = \"%s_%skb - %s_%skb l=%lld, g=%s, type=%s\" ]\n", t1.c_str(), t2.c_str(), t1.c_str(), s1.c_str(),
                t2.c_str(), s2.c_str(), s.links, avg.c_str(), s.type.c_str());
      }
    }
  }
  fprintf(gv, "}\n");
  fclose(gv);
}

// bin/LINKS.pl:653-658: upper-case, complement ATGCYRKMBDHV, reverse; anything else (N) stays
inline std::string reverse_complement(const std::string& s)
{
  static const char* from = "ATGCYRKMBDHV";
  static const char* to = "TACGRYMKVHDB";
  std::string out(s.size(), 'N');
  for (size_t i = 0; i < s.size(); ++i) {
    char ch = char(toupper((unsigned char)s[i]));
    if (const char* p = ch ? strchr(from, ch) : nullptr) ch = to[p - from];
    out[s.size() - 1 - i] = ch;
  }
  return out;
}

// bin/LINKS.pl:1306-1382
inline void build_scaffold_fasta(const std::vector<Scaffold>& scaffolds, const Contigs& tigs, const Options& o, const std::string& scaffold_file,
                                 const std::string& corr_file)
{
  const std::string fasta = scaffold_file + ".fa";
  FILE* out = fopen(fasta.c_str(), "w");
  if (!out) throw std::runtime_error("can't write to " + fasta + " -- fatal");
  FILE* corr = fopen(corr_file.c_str(), "w");
  if (!corr) { fclose(out); throw std::runtime_error("Can't write to " + corr_file + " -- fatal"); }
  fprintf(corr, "#This correspondence %s file lists the scaffold ID and contig ID used by LINKS (reported in the .scaffolds file), with the original contig name in the provided assembly file (-f %s)\n",
          corr_file.c_str(), o.assembly_file.c_str());
  fprintf(corr, "LINKS_scaffold_ID\tLINKS_contig_ID\toriginal_name\torientation(f=forward/r=reverse)\tnumber_of_links\tlinks_ratio\tgap_or_overlap(-)\n");
  for (const Scaffold& sc : scaffolds) {
    fprintf(out, ">scaffold%lld,%llu,%s\n", sc.number, (unsigned long long)sc.total, sc.text.c_str());
    for (const ChainStep& c : sc.chain) {
      const std::string name = size_t(c.tig) < tigs.name.size() ? tigs.name[size_t(c.tig)] : "";
      if (c.has_link)
        fprintf(corr, "scaffold%lld\t%lld\t%s\t%c\t%lld\t%s\t%lld\n", sc.number, c.tig, name.c_str(), c.orient, c.links, c.ratio.c_str(), c.mean);
      else
        fprintf(corr, "scaffold%lld\t%lld\t%s\t%c\tNA\tNA\tNA\n", sc.number, c.tig, name.c_str(), c.orient);
      const std::string& fwd = size_t(c.tig) < tigs.seq.size() ? tigs.seq[size_t(c.tig)] : std::string();
      if (c.orient == 'r') { const std::string rc = reverse_complement(fwd); fwrite(rc.data(), 1, rc.size(), out); }
      else fwrite(fwd.data(), 1, fwd.size(), out);
      if (c.has_link) {  // a gap is N x gap, an overlap (or 0) a single lower-case n (:1364-1365)
        if (c.mean > 0) { const std::string gap(size_t(c.mean), 'N'); fwrite(gap.data(), 1, gap.size(), out); }
        else fputc('n', out);
      }
    }
    fputc('\n', out);
  }
  fclose(corr);
  fclose(out);
}

inline std::string date_now()
{
  std::string d;
  if (FILE* p = popen("date", "r")) { read_line(p, d); pclose(p); }
  return d;
}

// The whole Perl stage (bin/LINKS.pl:117-186): reads <b>.tigpair_checkpoint.tsv, writes the scaffold files.
// rows != nullptr: the pairs come from memory (n_rows of them) and <b>.tigpair_checkpoint.tsv is not read.
inline int run(const Options& o, FILE* log_to = stdout, const RowIn* rows = nullptr, size_t n_rows = 0)
{
  const std::string b = o.base_name;
  const std::string scaffold = b + ".scaffolds", graph = b + ".gv", corr = b + ".assembly_correspondence.tsv";
  const std::string tigpair = b + ".tigpair_checkpoint.tsv", simplepair = b + ".simplepair_checkpoint.tsv", logf = b + ".log";
  FILE* lg = fopen(logf.c_str(), "a");
  auto say = [&](const std::string& m) { fputs(m.c_str(), log_to); if (lg) fputs(m.c_str(), lg); };
  if (!rows) {
    FILE* t = fopen(tigpair.c_str(), "r");
    if (!t) { fputs("No checkpoint provided", log_to); if (lg) fclose(lg); return 0; }  // :121-124
    fclose(t);
  }
  try {
    say("\n\nScaffolding mid-point files:\n" + tigpair + "\n" + simplepair + "\ndetected; LINKS will skip contig pairing and use info from files instead : " + date_now() + "\n");
    say("\nReading sequences from " + o.assembly_file + " : " + date_now() + "\n");
    const Contigs tigs = read_contigs_memory(o.assembly_file);
    say("\ndone.\nReading sequence pairs from " + tigpair + " : " + date_now() + "\n");
    LayoutState st;
    if (rows) pairs_from_rows(rows, n_rows, simplepair, tigs, o, st.pair, st.simple);
    else read_contig_pairs(tigpair, simplepair, tigs, o, st.pair, st.simple);
    say("\ndone.\nBuilding scaffolds : " + date_now() + "\n");
    const std::vector<Scaffold> scaffolds = build_scaffolds(st, tigs, o, scaffold);
    const std::string date = date_now();
    say("\ndone.\nBuilding .gv graph : " + date + "\n");
    build_graph(st.simple, tigs, o, graph, date);
    say("\n\n=>Scaffolding ended : " + date_now() + "\nScaffolds layout in: " + scaffold + "\nScaffold graph in: " + graph + "\n");
    say("\n\n=>Making FASTA file : " + date_now() + "\n");
    build_scaffold_fasta(scaffolds, tigs, o, scaffold, corr);
    say("Scaffolds FASTA in : " + scaffold + ".fa\n");
    say("\n\n=>Wrote correspondence file tracking LINKScontigID <=> OriginalContigNames : " + corr + "\n");
    say("\nScaffolding executed normally " + date_now() + "\n");
  } catch (const std::exception& e) {
    say(std::string("\nSomething went wrong running LINKS layout ") + date_now() + "\n" + e.what() + "\n");
    if (lg) fclose(lg);
    return 1;
  }
  if (lg) fclose(lg);
  return 0;
}

}  // namespace links_layout
