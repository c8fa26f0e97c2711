/* oracle/links_oracle_main.c -- command-line front end for the CPU oracle (TEST INFRASTRUCTURE).
 * usage: links_oracle -f contigs.fa -s reads.fof [-d d1,d2] [-k 15] [-t t1,t2] [-o 0] [-e 0.1]
 *                     [-p 0.001] -b base
 * Writes base.bloom, base.tigpair_checkpoint.tsv (sorted) and base.pairing_issues (sorted), the
 * three files LINKS_CPP writes (reference src/LINKS.hpp:328,808-809,1015-1040).
 */
#define _POSIX_C_SOURCE 200809L
#include "links_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

typedef struct { char* seq; uint64_t* off; uint64_t n, cap_seq, cap_off, len; } seqset;

static void ss_push_rec(seqset* s)
{
  if (s->n + 2 > s->cap_off) { s->cap_off = s->cap_off ? s->cap_off * 2 : 1024; s->off = realloc(s->off, s->cap_off * sizeof(uint64_t)); }
  if (s->n == 0) s->off[0] = 0;
  s->off[++s->n] = s->len;
}
static void ss_append(seqset* s, const char* p, size_t n)
{
  if (s->len + n + 1 > s->cap_seq) { while (s->len + n + 1 > s->cap_seq) s->cap_seq = s->cap_seq ? s->cap_seq * 2 : (1u << 20); s->seq = realloc(s->seq, s->cap_seq); }
  memcpy(s->seq + s->len, p, n);
  s->len += n;
}

/* FASTA (multi-line) / FASTQ (4-line) reader; records are appended to s. */
static int read_seqfile(const char* path, seqset* s)
{
  FILE* f = fopen(path, "r");
  if (!f) { fprintf(stderr, "cannot open %s\n", path); return -1; }
  char* line = NULL; size_t cap = 0; ssize_t n;
  int mode = 0; /* 0 none, 1 fasta seq, 2 fastq seq next, 3 fastq plus, 4 fastq qual */
  int open_rec = 0;
  while ((n = getline(&line, &cap, f)) >= 0) {
    while (n > 0 && (line[n - 1] == '\n' || line[n - 1] == '\r' || line[n - 1] == ' ' || line[n - 1] == '\t')) --n;
    if (mode == 2) { ss_append(s, line, (size_t)n); mode = 3; continue; }
    if (mode == 3) { mode = 4; continue; }
    if (mode == 4) { mode = 0; continue; }
    if (n == 0) continue;
    if (line[0] == '>') { if (open_rec) ss_push_rec(s); open_rec = 1; mode = 1; continue; }
    if (line[0] == '@' && mode != 1) { if (open_rec) ss_push_rec(s); open_rec = 1; mode = 2; continue; }
    if (mode == 1) ss_append(s, line, (size_t)n);
  }
  if (open_rec) ss_push_rec(s);
  free(line);
  fclose(f);
  return 0;
}

static unsigned parse_list(const char* txt, uint32_t* out, unsigned cap)
{
  unsigned n = 0;
  const char* p = txt;
  while (*p && n < cap) {
    out[n++] = (uint32_t)strtoul(p, NULL, 10);
    const char* c = strchr(p, ',');
    if (!c) break;
    p = c + 1;
  }
  return n;
}

int main(int argc, char** argv)
{
  const char *fa = NULL, *fof = NULL, *base = "oracle_out";
  uint32_t dist[256] = {4000}, stepl[256] = {2};
  unsigned n_dist = 1, n_step = 1, k = 15;
  uint64_t offset = 0;
  float e = 0.1f, fpr = 0.001f;
  int c;
  while ((c = getopt(argc, argv, "f:s:d:k:t:j:o:e:l:a:z:b:r:p:x:v:")) != -1) {
    switch (c) {
    case 'f': fa = optarg; break;
    case 's': fof = optarg; break;
    case 'd': n_dist = parse_list(optarg, dist, 256); break;
    case 't': n_step = parse_list(optarg, stepl, 256); break;
    case 'k': k = (unsigned)strtoul(optarg, NULL, 10); break;
    case 'o': offset = strtoull(optarg, NULL, 10); break;
    case 'e': e = strtof(optarg, NULL); break;
    case 'p': fpr = strtof(optarg, NULL); break;
    case 'b': base = optarg; break;
    default: break;
    }
  }
  if (!fa || !fof) { fprintf(stderr, "need -f and -s\n"); return 2; }
  uint16_t step[256];
  for (unsigned i = 0; i < n_dist; ++i) step[i] = (uint16_t)stepl[i < n_step ? i : n_step - 1];

  seqset contigs = {0}, reads = {0};
  if (read_seqfile(fa, &contigs)) return 1;
  FILE* ff = fopen(fof, "r");
  if (!ff) { fprintf(stderr, "cannot open %s\n", fof); return 1; }
  char path[4096];
  while (fscanf(ff, "%4095s", path) == 1)
    if (read_seqfile(path, &reads)) return 1;
  fclose(ff);

  struct stat st;
  if (stat(fa, &st)) return 1;
  uint64_t m_bits, bytes; unsigned hash_num;
  lo_bf_size((uint64_t)st.st_size, fpr, &m_bits, &bytes, &hash_num);
  printf("- m: %llu\n- hashFct: %u\n- bytes: %llu\n", (unsigned long long)m_bits, hash_num, (unsigned long long)bytes);
  uint8_t* bits = calloc(bytes, 1);
  for (uint64_t i = 0; i < contigs.n; ++i)
    lo_bf_insert_seq(bits, bytes, hash_num, k, contigs.seq + contigs.off[i], contigs.off[i + 1] - contigs.off[i]);

  char fn[4200], hdr[512];
  snprintf(fn, sizeof fn, "%s.bloom", base);
  FILE* o = fopen(fn, "wb");
  size_t hl = lo_bf_header(hdr, sizeof hdr, k, hash_num, bytes);
  fwrite(hdr, 1, hl, o); fwrite(bits, 1, bytes, o); fclose(o);

  lo_params p = {k, n_dist, dist, step, offset, e};
  uint64_t zero_off[1] = {0};
  lo_result* r = lo_run(&p, bits, bytes, hash_num, contigs.seq, contigs.n ? contigs.off : zero_off, contigs.n,
                        reads.seq, reads.n ? reads.off : zero_off, reads.n, 0);
  if (!r) return 1;
  char buf[1024];
  snprintf(fn, sizeof fn, "%s.tigpair_checkpoint.tsv", base);
  o = fopen(fn, "w");
  for (uint64_t i = 0; i < r->n_rows; ++i) { lo_format_row(&r->rows[i], buf, sizeof buf); fputs(buf, o); }
  fclose(o);
  snprintf(fn, sizeof fn, "%s.pairing_issues", base);
  o = fopen(fn, "w");
  for (uint64_t i = 0; i < r->n_issues; ++i) { lo_format_issue(&r->issues[i], buf, sizeof buf); fputs(buf, o); }
  fclose(o);
  printf("candidates %llu survivors %llu unique %llu mates %llu tracked %llu ok_pairs %llu problem %llu ok_contig %llu iz %llu illogical %llu gate_fail %llu rows %llu\n",
         (unsigned long long)r->stats.n_candidates, (unsigned long long)r->stats.n_survivors,
         (unsigned long long)r->stats.n_unique_pairs, (unsigned long long)r->stats.n_mates,
         (unsigned long long)r->stats.n_tracked, (unsigned long long)r->stats.ct_ok_pairs,
         (unsigned long long)r->stats.ct_problem_pairs, (unsigned long long)r->stats.ct_ok_contig,
         (unsigned long long)r->stats.ct_iz_issues, (unsigned long long)r->stats.ct_illogical,
         (unsigned long long)r->stats.ct_gate_fail, (unsigned long long)r->n_rows);
  lo_free(r);
  free(bits);
  return 0;
}
