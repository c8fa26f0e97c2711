// oracle/btl_probe.cpp -- TEST INFRASTRUCTURE ONLY.
//
// A thin extern "C" shim that #includes the *reference's own* btllib headers from
// /root/reference/btllib/include (nothing is copied) so the tests can call the real
// btllib::NtHash / btllib::KmerBloomFilter on arbitrary inputs and compare them with the
// plain-C restatement in links_oracle.c.  Built only where /root/reference exists
// (`make -C oracle ref`), into oracle/_ref/libbtl_probe.so.
#include "btllib/bloom_filter.hpp"
#include "btllib/nthash.hpp"
#include "btllib/seq_reader.hpp"

#include <cstdint>
#include <cstring>
#include <string>

extern "C" {

// Walk btllib::NtHash over seq from `start`; record for every successful roll() the position
// and fwd/rev hashes and all `hash_num` multi-hashes.  Returns the number of rolls written
// (<= cap).  (btllib/include/btllib/nthash.hpp:1265-1338,1493-1550)
uint64_t
btl_nthash_walk(const char* seq, uint64_t len, unsigned hash_num, unsigned k, uint64_t start,
                uint64_t cap, uint64_t* pos, uint64_t* fwd, uint64_t* rev, uint64_t* hashes)
{
  btllib::NtHash nt(seq, len, hash_num, k, start);
  uint64_t n = 0;
  while (n < cap && nt.roll()) {
    pos[n] = nt.get_pos();
    fwd[n] = nt.get_forward_hash();
    rev[n] = nt.get_reverse_hash();
    if (hashes) std::memcpy(hashes + n * hash_num, nt.hashes(), sizeof(uint64_t) * hash_num);
    ++n;
  }
  return n;
}

// Build a btllib::KmerBloomFilter of `bytes` bytes / hash_num / k, insert every sequence, and
// copy the raw bit array to `out` (must hold the rounded byte count, returned).
// (btllib/include/btllib/bloom_filter.hpp:617-646,807-821)
uint64_t
btl_bf_build(uint64_t bytes, unsigned hash_num, unsigned k, const char* seqs, const uint64_t* off,
             uint64_t n_seq, uint8_t* out, const char* save_path)
{
  btllib::KmerBloomFilter bf(bytes, hash_num, k);
  for (uint64_t i = 0; i < n_seq; ++i) bf.insert(seqs + off[i], off[i + 1] - off[i]);
  const uint64_t nb = bf.get_bytes();
  if (save_path && save_path[0]) bf.save(save_path);
  if (out) {
    // the bit array is private; query bit by bit through the public API instead
    btllib::BloomFilter& raw = bf.get_bloom_filter();
    for (uint64_t bit = 0; bit < nb * 8; ++bit) {
      // hashes[i] % array_bits == bit for hashes[i] = bit (bit < array_bits)
      uint64_t h[1024];
      for (unsigned j = 0; j < hash_num; ++j) h[j] = bit;
      if (raw.contains(h)) out[bit >> 3] |= uint8_t(1u << (bit & 7));
    }
  }
  return nb;
}

// Read `path` with btllib::SeqReader exactly as LINKS opens its inputs (Flag::LONG_MODE, src/LINKS.hpp:415-419); for
// record i (i < cap): lens[i] = bases, fnv[i] = FNV-1a of the sequence, nums[i] = record.num.  Returns the record count.
uint64_t
btl_read_records(const char* path, uint64_t cap, uint64_t* lens, uint64_t* fnv, uint64_t* nums)
{
  btllib::SeqReader reader(path, btllib::SeqReader::Flag::LONG_MODE);
  uint64_t n = 0;
  for (auto record : reader) {
    if (n < cap) {
      uint64_t h = 1469598103934665603ull;
      for (unsigned char c : record.seq) h = (h ^ c) * 1099511628211ull;
      lens[n] = record.seq.size(); fnv[n] = h; nums[n] = record.num;
    }
    ++n;
  }
  return n;
}

} // extern "C"
