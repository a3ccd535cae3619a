/*
 * snappy_lite.h -- the snappy RAW block format (what ovFile frames its .ovb blocks with: ST/ovStoreFile.C:157-190 calls
 * snappy::RawCompress / RawUncompress of the vendored ST/libsnappy).  Written from the format description
 * (varint length, then literal / copy elements); any valid stream is accepted by the stock decoder, so the compressor
 * need not reproduce the reference's byte stream -- only the decompressed words matter.
 *
 *   element tags (low two bits of the tag byte)
 *     00 literal   length-1 in the upper six bits when < 60, else 60..63 = 1..4 following length bytes (little endian)
 *     01 copy      length 4..11 in bits 2-4 (minus 4), offset 0..2047: bits 5-7 = offset >> 8, next byte = offset & 255
 *     10 copy      length 1..64 in the upper six bits (minus 1), next two bytes = offset (little endian)
 *     11 copy      as 10 with a four-byte offset
 */
#ifndef OIC_HOST_SNAPPY_LITE_H
#define OIC_HOST_SNAPPY_LITE_H
#include <stdint.h>
#include <string.h>

#include <vector>

namespace oichost {

inline size_t snappy_max_compressed(size_t n) { return 32 + n + n / 6; }

namespace snappy_detail {
inline uint32_t load32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline uint8_t *emit_literal(uint8_t *op, const uint8_t *lit, size_t len) {
  if (len == 0) return op;
  const size_t n = len - 1;
  if (n < 60) *op++ = (uint8_t)(n << 2);
  else if (n < (1u << 8)) { *op++ = 60 << 2; *op++ = (uint8_t)n; }
  else if (n < (1u << 16)) { *op++ = 61 << 2; *op++ = (uint8_t)n; *op++ = (uint8_t)(n >> 8); }
  else if (n < (1u << 24)) { *op++ = 62 << 2; *op++ = (uint8_t)n; *op++ = (uint8_t)(n >> 8); *op++ = (uint8_t)(n >> 16); }
  else { *op++ = 63 << 2; *op++ = (uint8_t)n; *op++ = (uint8_t)(n >> 8); *op++ = (uint8_t)(n >> 16); *op++ = (uint8_t)(n >> 24); }
  memcpy(op, lit, len);
  return op + len;
}
inline uint8_t *emit_copy_upto64(uint8_t *op, size_t offset, size_t l) {   // 4 <= l <= 64, offset <= 65535
  if (l <= 11 && offset < 2048) {
    *op++ = (uint8_t)(1 | ((l - 4) << 2) | ((offset >> 8) << 5));
    *op++ = (uint8_t)offset;
  } else {
    *op++ = (uint8_t)(2 | ((l - 1) << 2));
    *op++ = (uint8_t)offset;
    *op++ = (uint8_t)(offset >> 8);
  }
  return op;
}
inline uint8_t *emit_copy(uint8_t *op, size_t offset, size_t len) {   // len >= 4; pieces of at most 64, never a tail shorter than 4
  while (len >= 68) { op = emit_copy_upto64(op, offset, 64); len -= 64; }
  if (len > 64) { op = emit_copy_upto64(op, offset, 60); len -= 60; }
  return emit_copy_upto64(op, offset, len);
}
}  // namespace snappy_detail

/* greedy single-pass LZ77 with a 2^14-entry hash of 4-byte sequences; out must hold snappy_max_compressed(n) bytes */
inline size_t snappy_compress(const uint8_t *in, size_t n, uint8_t *out) {
  using namespace snappy_detail;
  uint8_t *op = out;
  size_t v = n;
  while (v >= 0x80) { *op++ = (uint8_t)(v | 0x80); v >>= 7; }
  *op++ = (uint8_t)v;
  if (n < 16) return (size_t)(emit_literal(op, in, n) - out);
  std::vector<uint32_t> table(1u << 14, 0xffffffffu);
  size_t ip = 0, lit = 0;
  const size_t limit = n - 4;
  while (ip <= limit) {
    const uint32_t cur = load32(in + ip);
    const uint32_t h = (cur * 0x1e35a7bdu) >> 18;
    const uint32_t cand = table[h];
    table[h] = (uint32_t)ip;
    if (cand != 0xffffffffu && ip - cand <= 65535 && load32(in + cand) == cur) {
      size_t len = 4;
      while (ip + len < n && in[cand + len] == in[ip + len]) len++;
      op = emit_literal(op, in + lit, ip - lit);
      op = emit_copy(op, ip - cand, len);
      ip += len;
      lit = ip;
    } else ip++;
  }
  op = emit_literal(op, in + lit, n - lit);
  return (size_t)(op - out);
}

/* returns false on a malformed stream */
inline bool snappy_uncompress(const uint8_t *in, size_t n, std::vector<uint8_t> &out) {
  size_t ip = 0, total = 0;
  int shift = 0;
  while (true) {
    if (ip >= n || shift > 35) return false;
    const uint8_t b = in[ip++];
    total |= (size_t)(b & 0x7f) << shift;
    if (!(b & 0x80)) break;
    shift += 7;
  }
  out.resize(total);
  size_t op = 0;
  while (ip < n) {
    const uint8_t tag = in[ip++];
    size_t len, off;
    switch (tag & 3) {
      case 0: {
        len = (tag >> 2) + 1;
        if (len > 60) {
          const size_t nb = len - 60;
          if (ip + nb > n) return false;
          len = 0;
          for (size_t k = 0; k < nb; k++) len |= (size_t)in[ip + k] << (8 * k);
          len += 1;
          ip += nb;
        }
        if (ip + len > n || op + len > total) return false;
        memcpy(out.data() + op, in + ip, len);
        ip += len; op += len;
        continue;
      }
      case 1:
        if (ip + 1 > n) return false;
        len = ((tag >> 2) & 7) + 4;
        off = ((size_t)(tag >> 5) << 8) | in[ip];
        ip += 1;
        break;
      case 2:
        if (ip + 2 > n) return false;
        len = (tag >> 2) + 1;
        off = (size_t)in[ip] | ((size_t)in[ip + 1] << 8);
        ip += 2;
        break;
      default:
        if (ip + 4 > n) return false;
        len = (tag >> 2) + 1;
        off = (size_t)in[ip] | ((size_t)in[ip + 1] << 8) | ((size_t)in[ip + 2] << 16) | ((size_t)in[ip + 3] << 24);
        ip += 4;
        break;
    }
    if (off == 0 || off > op || op + len > total) return false;
    for (size_t k = 0; k < len; k++) out[op + k] = out[op + k - off];   // byte by byte: copies may overlap their own output
    op += len;
  }
  return op == total;
}

}  // namespace oichost
#endif
