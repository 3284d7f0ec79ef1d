// Host-side (CPU) crypto for 16..1024-byte control values only: BLAKE3 for
// SessionId / DeriveSeed and the AES-128 key schedule.  No tensor data is ever
// processed here -- tensor-sized PRF work happens in sample.cu / rep.cu on the GPU.
//
// DeriveSeed is one hash per sampled tensor, so it stays on
// the CPU exactly like in the reference (moose/src/host/prim.rs:123-149).
#include "common.h"

#include <sys/random.h>

namespace {

// ---- BLAKE3 (single chunk, inputs <= 1024 bytes) ----------------------------
const uint32_t B3_IV[8] = {0x6A09E667u, 0xBB67AE85u, 0x3C6EF372u, 0xA54FF53Au,
                           0x510E527Fu, 0x9B05688Cu, 0x1F83D9ABu, 0x5BE0CD19u};
const int B3_PERM[16] = {2, 6, 3, 10, 7, 0, 4, 13, 1, 11, 12, 5, 9, 14, 15, 8};
enum { B3_CHUNK_START = 1, B3_CHUNK_END = 2, B3_ROOT = 8, B3_KEYED = 16, B3_DK_CONTEXT = 32, B3_DK_MATERIAL = 64 };

inline uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

inline void g(uint32_t *s, int a, int b, int c, int d, uint32_t mx, uint32_t my) {
  s[a] = s[a] + s[b] + mx; s[d] = rotr(s[d] ^ s[a], 16);
  s[c] = s[c] + s[d];      s[b] = rotr(s[b] ^ s[c], 12);
  s[a] = s[a] + s[b] + my; s[d] = rotr(s[d] ^ s[a], 8);
  s[c] = s[c] + s[d];      s[b] = rotr(s[b] ^ s[c], 7);
}

void compress(const uint32_t cv[8], const uint8_t block[64], uint8_t block_len, uint64_t counter, uint32_t flags,
              uint32_t out[16]) {
  uint32_t m[16];
  for (int i = 0; i < 16; i++)
    m[i] = (uint32_t)block[4 * i] | ((uint32_t)block[4 * i + 1] << 8) | ((uint32_t)block[4 * i + 2] << 16) |
           ((uint32_t)block[4 * i + 3] << 24);
  uint32_t s[16] = {cv[0], cv[1], cv[2], cv[3], cv[4], cv[5], cv[6], cv[7],
                    B3_IV[0], B3_IV[1], B3_IV[2], B3_IV[3],
                    (uint32_t)counter, (uint32_t)(counter >> 32), block_len, flags};
  for (int r = 0; r < 7; r++) {
    g(s, 0, 4, 8, 12, m[0], m[1]);
    g(s, 1, 5, 9, 13, m[2], m[3]);
    g(s, 2, 6, 10, 14, m[4], m[5]);
    g(s, 3, 7, 11, 15, m[6], m[7]);
    g(s, 0, 5, 10, 15, m[8], m[9]);
    g(s, 1, 6, 11, 12, m[10], m[11]);
    g(s, 2, 7, 8, 13, m[12], m[13]);
    g(s, 3, 4, 9, 14, m[14], m[15]);
    if (r < 6) {
      uint32_t p[16];
      for (int i = 0; i < 16; i++) p[i] = m[B3_PERM[i]];
      memcpy(m, p, sizeof(m));
    }
  }
  for (int i = 0; i < 8; i++) {
    out[i] = s[i] ^ s[i + 8];
    out[i + 8] = s[i + 8] ^ cv[i];
  }
}

// Hash of one chunk (len <= 1024); writes the first out_len (<= 64) bytes of the root output.
void hash_chunk(const uint32_t key[8], uint32_t mode, const uint8_t *in, size_t len, uint8_t *out, size_t out_len) {
  uint32_t cv[8];
  memcpy(cv, key, sizeof(cv));
  size_t nblocks = len == 0 ? 1 : (len + 63) / 64;
  uint32_t words[16];
  for (size_t b = 0; b < nblocks; b++) {
    uint8_t block[64] = {0};
    size_t bl = (b + 1 == nblocks) ? len - 64 * b : 64;
    if (bl) memcpy(block, in + 64 * b, bl);
    uint32_t flags = mode;
    if (b == 0) flags |= B3_CHUNK_START;
    if (b + 1 == nblocks) flags |= B3_CHUNK_END | B3_ROOT;
    compress(cv, block, (uint8_t)bl, 0, flags, words);
    memcpy(cv, words, sizeof(cv));
  }
  uint8_t full[64];
  for (int i = 0; i < 16; i++) {
    full[4 * i] = (uint8_t)words[i];
    full[4 * i + 1] = (uint8_t)(words[i] >> 8);
    full[4 * i + 2] = (uint8_t)(words[i] >> 16);
    full[4 * i + 3] = (uint8_t)(words[i] >> 24);
  }
  memcpy(out, full, out_len);
}

void key_words(const uint8_t k[32], uint32_t w[8]) {
  for (int i = 0; i < 8; i++)
    w[i] = (uint32_t)k[4 * i] | ((uint32_t)k[4 * i + 1] << 8) | ((uint32_t)k[4 * i + 2] << 16) |
           ((uint32_t)k[4 * i + 3] << 24);
}

// ---- AES-128 key schedule / single block (control values only) --------------
const uint8_t SBOX[256] = {
    0x63, 0x7c, 0x77, 0x7b, 0xf2, 0x6b, 0x6f, 0xc5, 0x30, 0x01, 0x67, 0x2b, 0xfe, 0xd7, 0xab, 0x76, 0xca, 0x82, 0xc9,
    0x7d, 0xfa, 0x59, 0x47, 0xf0, 0xad, 0xd4, 0xa2, 0xaf, 0x9c, 0xa4, 0x72, 0xc0, 0xb7, 0xfd, 0x93, 0x26, 0x36, 0x3f,
    0xf7, 0xcc, 0x34, 0xa5, 0xe5, 0xf1, 0x71, 0xd8, 0x31, 0x15, 0x04, 0xc7, 0x23, 0xc3, 0x18, 0x96, 0x05, 0x9a, 0x07,
    0x12, 0x80, 0xe2, 0xeb, 0x27, 0xb2, 0x75, 0x09, 0x83, 0x2c, 0x1a, 0x1b, 0x6e, 0x5a, 0xa0, 0x52, 0x3b, 0xd6, 0xb3,
    0x29, 0xe3, 0x2f, 0x84, 0x53, 0xd1, 0x00, 0xed, 0x20, 0xfc, 0xb1, 0x5b, 0x6a, 0xcb, 0xbe, 0x39, 0x4a, 0x4c, 0x58,
    0xcf, 0xd0, 0xef, 0xaa, 0xfb, 0x43, 0x4d, 0x33, 0x85, 0x45, 0xf9, 0x02, 0x7f, 0x50, 0x3c, 0x9f, 0xa8, 0x51, 0xa3,
    0x40, 0x8f, 0x92, 0x9d, 0x38, 0xf5, 0xbc, 0xb6, 0xda, 0x21, 0x10, 0xff, 0xf3, 0xd2, 0xcd, 0x0c, 0x13, 0xec, 0x5f,
    0x97, 0x44, 0x17, 0xc4, 0xa7, 0x7e, 0x3d, 0x64, 0x5d, 0x19, 0x73, 0x60, 0x81, 0x4f, 0xdc, 0x22, 0x2a, 0x90, 0x88,
    0x46, 0xee, 0xb8, 0x14, 0xde, 0x5e, 0x0b, 0xdb, 0xe0, 0x32, 0x3a, 0x0a, 0x49, 0x06, 0x24, 0x5c, 0xc2, 0xd3, 0xac,
    0x62, 0x91, 0x95, 0xe4, 0x79, 0xe7, 0xc8, 0x37, 0x6d, 0x8d, 0xd5, 0x4e, 0xa9, 0x6c, 0x56, 0xf4, 0xea, 0x65, 0x7a,
    0xae, 0x08, 0xba, 0x78, 0x25, 0x2e, 0x1c, 0xa6, 0xb4, 0xc6, 0xe8, 0xdd, 0x74, 0x1f, 0x4b, 0xbd, 0x8b, 0x8a, 0x70,
    0x3e, 0xb5, 0x66, 0x48, 0x03, 0xf6, 0x0e, 0x61, 0x35, 0x57, 0xb9, 0x86, 0xc1, 0x1d, 0x9e, 0xe1, 0xf8, 0x98, 0x11,
    0x69, 0xd9, 0x8e, 0x94, 0x9b, 0x1e, 0x87, 0xe9, 0xce, 0x55, 0x28, 0xdf, 0x8c, 0xa1, 0x89, 0x0d, 0xbf, 0xe6, 0x42,
    0x68, 0x41, 0x99, 0x2d, 0x0f, 0xb0, 0x54, 0xbb, 0x16};

inline uint32_t sub_word(uint32_t w) {
  return (uint32_t)SBOX[w & 0xff] | ((uint32_t)SBOX[(w >> 8) & 0xff] << 8) | ((uint32_t)SBOX[(w >> 16) & 0xff] << 16) |
         ((uint32_t)SBOX[w >> 24] << 24);
}
inline uint8_t xt(uint8_t x) { return (uint8_t)((x << 1) ^ ((x >> 7) * 0x1b)); }

}  // namespace

void mb_blake3_session_id(const char *s, size_t len, uint8_t out[16]) {
  // blake3::Hasher::new().update(s).finalize_xof().fill(16 bytes)  (computation.rs:108-124)
  if (len > 1024) len = 1024;  // guarded by the caller
  hash_chunk(B3_IV, 0, (const uint8_t *)s, len, out, 16);
}

void mb_blake3_derive_seed(const uint8_t sid[16], const uint8_t sync_key[16], const uint8_t key[16], uint8_t out[16]) {
  // derive_key("Derive Seed", key) -> keyed hash over sid || sync_key -> first 16 XOF bytes
  static const char ctxstr[] = "Derive Seed";
  uint8_t context_key[32], derived[32];
  uint32_t w[8];
  hash_chunk(B3_IV, B3_DK_CONTEXT, (const uint8_t *)ctxstr, sizeof(ctxstr) - 1, context_key, 32);
  key_words(context_key, w);
  hash_chunk(w, B3_DK_MATERIAL, key, 16, derived, 32);
  key_words(derived, w);
  uint8_t nonce[32];
  memcpy(nonce, sid, 16);
  memcpy(nonce + 16, sync_key, 16);
  hash_chunk(w, B3_KEYED, nonce, 32, out, 16);
}

void mb_aes_expand(const uint8_t key[16], AesKey *out) {
  static const uint8_t rcon[10] = {1, 2, 4, 8, 16, 32, 64, 128, 0x1b, 0x36};
  uint32_t *w = out->rk;
  for (int i = 0; i < 4; i++)
    w[i] = (uint32_t)key[4 * i] | ((uint32_t)key[4 * i + 1] << 8) | ((uint32_t)key[4 * i + 2] << 16) |
           ((uint32_t)key[4 * i + 3] << 24);
  for (int i = 4; i < 44; i++) {
    uint32_t t = w[i - 1];
    if (i % 4 == 0) {
      t = (t >> 8) | (t << 24);  // RotWord on little-endian column words
      t = sub_word(t) ^ rcon[i / 4 - 1];
    }
    w[i] = w[i - 4] ^ t;
  }
  // Round-1 constants for counter blocks LE128(ctr), ctr < 2^32: state words 1..3 are rk[1..3].
  auto te0 = [](uint32_t x) {
    uint32_t sb = SBOX[x & 0xff];
    uint32_t s2 = xt((uint8_t)sb), s3 = s2 ^ sb;
    return s2 | (sb << 8) | (sb << 16) | (s3 << 24);
  };
  auto rotl = [](uint32_t v, int n) { return (v << n) | (v >> (32 - n)); };
  auto col = [&](uint32_t b0, int use0, uint32_t b1, int use1, uint32_t b2, int use2, uint32_t b3, int use3, uint32_t rk) {
    uint32_t v = rk;
    if (use0) v ^= te0(b0);
    if (use1) v ^= rotl(te0(b1 >> 8), 8);
    if (use2) v ^= rotl(te0(b2 >> 16), 16);
    if (use3) v ^= rotl(te0(b3 >> 24), 24);
    return v;
  };
  const uint32_t s1 = w[1], s2 = w[2], s3 = w[3];
  out->r1c[0] = col(0, 0, s1, 1, s2, 1, s3, 1, w[4]);  // t0 = Te0[s0.b0] ^ ...
  out->r1c[1] = col(s1, 1, s2, 1, s3, 1, 0, 0, w[5]);  // t1 = ... ^ rotl24(Te0[s0.b3])
  out->r1c[2] = col(s2, 1, s3, 1, 0, 0, s1, 1, w[6]);  // t2 = ... ^ rotl16(Te0[s0.b2]) ...
  out->r1c[3] = col(s3, 1, 0, 0, s1, 1, s2, 1, w[7]);  // t3 = ... ^ rotl8(Te0[s0.b1]) ...
}

void mb_aes_encrypt_host(const AesKey *k, const uint8_t in[16], uint8_t out[16]) {
  uint8_t s[16], t[16];
  const uint8_t *rk = (const uint8_t *)k->rk;  // little-endian host: words == byte stream
  for (int i = 0; i < 16; i++) s[i] = in[i] ^ rk[i];
  for (int r = 1; r <= 10; r++) {
    for (int c = 0; c < 4; c++)
      for (int row = 0; row < 4; row++) t[4 * c + row] = SBOX[s[4 * ((c + row) & 3) + row]];
    if (r < 10) {
      for (int c = 0; c < 4; c++) {
        uint8_t a0 = t[4 * c], a1 = t[4 * c + 1], a2 = t[4 * c + 2], a3 = t[4 * c + 3];
        uint8_t x = a0 ^ a1 ^ a2 ^ a3;
        s[4 * c] = a0 ^ x ^ xt(a0 ^ a1);
        s[4 * c + 1] = a1 ^ x ^ xt(a1 ^ a2);
        s[4 * c + 2] = a2 ^ x ^ xt(a2 ^ a3);
        s[4 * c + 3] = a3 ^ x ^ xt(a3 ^ a0);
      }
    } else {
      memcpy(s, t, 16);
    }
    for (int i = 0; i < 16; i++) s[i] ^= rk[16 * r + i];
  }
  memcpy(out, s, 16);
}

int mb_entropy16(mb_ctx *ctx, uint8_t out[16]) {
  std::lock_guard<std::mutex> lock(ctx->mu);
  if (ctx->deterministic) {
    AesKey k;
    mb_aes_expand(ctx->entropy_key, &k);
    uint8_t ctr[16] = {0};
    uint64_t c = ctx->entropy_counter++;
    memcpy(ctr, &c, 8);
    mb_aes_encrypt_host(&k, ctr, out);
    return MB_OK;
  }
  size_t got = 0;
  while (got < 16) {
    ssize_t r = getrandom(out + got, 16 - got, 0);
    if (r < 0) return mb_fail(ctx, MB_ERR_KERNEL, "getrandom failed");
    got += (size_t)r;
  }
  return MB_OK;
}

// Context-free self-test hooks (CPU only; used by the `-m "not gpu"` tests to check
// the host-side hashing / key schedule without a device).
extern "C" void mb_selftest_session_id(const char *s, uint8_t out[16]) { mb_blake3_session_id(s, strlen(s), out); }
extern "C" void mb_selftest_derive_seed(const uint8_t sid[16], const uint8_t sync_key[16], const uint8_t key[16],
                                        uint8_t out[16]) {
  mb_blake3_derive_seed(sid, sync_key, key, out);
}
extern "C" void mb_selftest_aes_block(const uint8_t key[16], const uint8_t in[16], uint8_t out[16]) {
  AesKey k;
  mb_aes_expand(key, &k);
  mb_aes_encrypt_host(&k, in, out);
}
