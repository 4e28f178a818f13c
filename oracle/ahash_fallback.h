/* ORACLE (test infrastructure only — never linked into or called by the product path).
 *
 * CPU restatement of ahash 0.8.x `fallback_hash::AHasher` as used by DataFusion v34's
 * `create_hashes` through `RandomState::with_seeds(0,0,0,0)` (BatchPartitioner, "use fixed random
 * state"). The algorithm lives in the un-vendored crate `ahash = "0.8"` (reference pin:
 * /root/reference/ballista/core/Cargo.toml:47; no Cargo.lock). Reference call site:
 * /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-346.
 *
 * Structure pinned against compiled ahash-0.8.12 machine code (oracle/evidence/README.md):
 *   buffer = k1, pad = k0, extra_keys = [k2, k3], k_i = seed_i ^ PI2[i].
 * and against known answers produced by EXECUTING compiled ahash-0.8.12 (tests/golden/ahash_0_8_12_kat.json:
 * hash_one::<&[u8]> and from_keys, called by address in a Rust cdylib of this image; tests/test_oracle_ahash_kat.py).
 * Still recalled, not executed: with_seeds' k_i = seed_i ^ PI2[i]. No golden value from a REFERENCE build exists on this
 * box (no Rust), so DataFusion's glue around the hasher stays "parity unpinned".
 */
#ifndef ORACLE_AHASH_FALLBACK_H
#define ORACLE_AHASH_FALLBACK_H
#include <stdint.h>
#include <string.h>

#define AH_MULTIPLE 6364136223846793005ULL /* ahash/src/operations.rs MULTIPLE */
#define AH_ROT 23                          /* ahash/src/fallback_hash.rs ROT */

static const uint64_t AH_PI2[4] = {/* ahash/src/random_state.rs PI2 */
                                   0x452821e638d01377ULL, 0xbe5466cf34e90c6cULL,
                                   0xc0ac29b7c97c50ddULL, 0x3f84d5b5b5470917ULL};

typedef struct {
  uint64_t buffer, pad, extra0, extra1;
} ah_hasher;

static inline uint64_t ah_rotl(uint64_t x, unsigned r) {
  r &= 63;
  return r ? (x << r) | (x >> (64 - r)) : x;
}

/* operations.rs folded_multiply (64-bit targets: full 128-bit product, lo ^ hi) */
static inline uint64_t ah_folded_multiply(uint64_t s, uint64_t by) {
  unsigned __int128 r = (unsigned __int128)s * (unsigned __int128)by;
  return (uint64_t)r ^ (uint64_t)(r >> 64);
}

/* RandomState::with_seeds(s0..s3) followed by AHasher::from_random_state */
static inline ah_hasher ah_new(uint64_t s0, uint64_t s1, uint64_t s2, uint64_t s3) {
  ah_hasher h;
  uint64_t k0 = s0 ^ AH_PI2[0], k1 = s1 ^ AH_PI2[1], k2 = s2 ^ AH_PI2[2], k3 = s3 ^ AH_PI2[3];
  h.buffer = k1;
  h.pad = k0;
  h.extra0 = k2;
  h.extra1 = k3;
  return h;
}

/* write_u8/u16/u32/u64: update(i as u64) */
static inline void ah_update(ah_hasher* h, uint64_t x) {
  h->buffer = ah_folded_multiply(x ^ h->buffer, AH_MULTIPLE);
}

/* write_u128 and the body of write(): large_update */
static inline void ah_large_update(ah_hasher* h, uint64_t lo, uint64_t hi) {
  uint64_t combined = ah_folded_multiply(lo ^ h->extra0, hi ^ h->extra1);
  h->buffer = ah_rotl((h->buffer + h->pad) ^ combined, AH_ROT);
}

static inline uint64_t ah_rd(const uint8_t* p, int n) { /* little-endian read of n<=8 bytes */
  uint64_t v = 0;
  memcpy(&v, p, (size_t)n);
  return v;
}

/* Hasher::write(bytes) */
static inline void ah_write(ah_hasher* h, const uint8_t* d, uint64_t len) {
  h->buffer = (h->buffer + len) * AH_MULTIPLE;
  if (len > 8) {
    if (len > 16) {
      ah_large_update(h, ah_rd(d + len - 16, 8), ah_rd(d + len - 8, 8));
      while (len > 16) {
        ah_large_update(h, ah_rd(d, 8), ah_rd(d + 8, 8));
        d += 16;
        len -= 16;
      }
    } else {
      ah_large_update(h, ah_rd(d, 8), ah_rd(d + len - 8, 8));
    }
  } else if (len >= 4) {
    ah_large_update(h, ah_rd(d, 4), ah_rd(d + len - 4, 4));
  } else if (len >= 2) {
    ah_large_update(h, ah_rd(d, 2), (uint64_t)d[len - 1]);
  } else if (len == 1) {
    ah_large_update(h, (uint64_t)d[0], (uint64_t)d[0]);
  } else {
    ah_large_update(h, 0, 0);
  }
}

static inline uint64_t ah_finish(const ah_hasher* h) {
  return ah_rotl(ah_folded_multiply(h->buffer, h->pad), (unsigned)(h->buffer & 63));
}

/* hash_one for the Rust types DataFusion feeds (SURVEY.md A.1 "What Rust's Hash feeds") */
static inline uint64_t ah_hash_u64(uint64_t x) { /* i8..i64,u8..u64 zero-extended from own width; bool */
  ah_hasher h = ah_new(0, 0, 0, 0);
  ah_update(&h, x);
  return ah_finish(&h);
}
static inline uint64_t ah_hash_u128(uint64_t lo, uint64_t hi) { /* i128 / Decimal128 */
  ah_hasher h = ah_new(0, 0, 0, 0);
  ah_large_update(&h, lo, hi);
  return ah_finish(&h);
}
static inline uint64_t ah_hash_str(const uint8_t* d, uint64_t len) { /* str: write(bytes); write_u8(0xff) */
  ah_hasher h = ah_new(0, 0, 0, 0);
  ah_write(&h, d, len);
  ah_update(&h, 0xff);
  return ah_finish(&h);
}
static inline uint64_t ah_hash_bytes(const uint8_t* d, uint64_t len) { /* [u8]: write_usize(len); write */
  ah_hasher h = ah_new(0, 0, 0, 0);
  ah_update(&h, len);
  ah_write(&h, d, len);
  return ah_finish(&h);
}
/* datafusion-common hash_utils.rs combine_hashes(l, r) = (17*37 + l)*37 + r, wrapping */
static inline uint64_t df_combine_hashes(uint64_t l, uint64_t r) {
  return (17ULL * 37ULL + l) * 37ULL + r;
}
#endif
