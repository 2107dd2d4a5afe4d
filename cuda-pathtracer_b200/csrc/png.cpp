// png.cpp -- ptc_write_png: the RGBA8 frame as a PNG file, host only, no third-party code.
//
// the reference shows its frame in a window (interop.cpp:67-72 blits the GL
// surface raytrace.cu:260-270 wrote); headless, the observable result is a file.  A PNG is an
// 8-byte signature, IHDR, the zlib stream of the filtered scanlines in IDAT, IEND (RFC 2083).
// The zlib stream here is one deflate block with the FIXED Huffman code (RFC 1951 3.2.6) over a
// greedy LZ77 parse (one hash probe per position, 32 KB window): a path-traced frame is noise
// that hardly compresses, the flat regions of an early / preview frame do.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "ptcore.h"

namespace {

struct BitWriter
{
  std::vector<uint8_t>& out;
  uint64_t acc = 0;
  int n = 0;
  explicit BitWriter(std::vector<uint8_t>& o) : out(o) {}
  void put(uint32_t bits, int count) // least significant bit first (RFC 1951 3.1.1)
  {
    acc |= (uint64_t)bits << n;
    n += count;
    while (n >= 8) {
      out.push_back((uint8_t)acc);
      acc >>= 8;
      n -= 8;
    }
  }
  void put_code(uint32_t code, int count) // Huffman codes go most significant bit first
  {
    uint32_t r = 0;
    for (int i = 0; i < count; ++i)
      r |= ((code >> i) & 1u) << (count - 1 - i);
    put(r, count);
  }
  void flush()
  {
    if (n)
      put(0, 8 - n);
  }
};

void
put_literal_or_length_symbol(BitWriter& w, unsigned sym)
{
  if (sym < 144)
    w.put_code(0x30 + sym, 8);
  else if (sym < 256)
    w.put_code(0x190 + (sym - 144), 9);
  else if (sym < 280)
    w.put_code(sym - 256, 7);
  else
    w.put_code(0xC0 + (sym - 280), 8);
}

const uint16_t kLenBase[29] = { 3,  4,  5,  6,  7,  8,  9,  10, 11,  13,  15,  17,  19,  23, 27,
                                31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258 };
const uint8_t kLenExtra[29] = { 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0 };
const uint16_t kDistBase[30] = { 1,   2,   3,   4,   5,   7,    9,    13,   17,   25,   33,   49,   65,    97,    129,
                                 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577 };
const uint8_t kDistExtra[30] = { 0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13 };

void
put_match(BitWriter& w, unsigned len, unsigned dist)
{
  int ls = 28;
  while (kLenBase[ls] > len)
    --ls;
  put_literal_or_length_symbol(w, 257 + ls);
  w.put(len - kLenBase[ls], kLenExtra[ls]);
  int ds = 29;
  while (kDistBase[ds] > dist)
    --ds;
  w.put_code(ds, 5);
  w.put(dist - kDistBase[ds], kDistExtra[ds]);
}

// zlib stream (RFC 1950) of `n` bytes
void
zlib_compress(const uint8_t* src, size_t n, std::vector<uint8_t>& out)
{
  out.push_back(0x78); // deflate, 32 KB window
  out.push_back(0x01); // no preset dictionary, check bits: 0x7801 % 31 == 0
  BitWriter w(out);
  w.put(1, 1); // last block
  w.put(1, 2); // fixed Huffman code
  constexpr int kHashBits = 15;
  std::vector<int64_t> head((size_t)1 << kHashBits, -1);
  auto hash3 = [&](size_t i) {
    const uint32_t v = src[i] | (src[i + 1] << 8) | (src[i + 2] << 16);
    return (v * 2654435761u) >> (32 - kHashBits);
  };
  size_t i = 0;
  while (i < n) {
    unsigned best = 0, dist = 0;
    if (i + 3 <= n) {
      const uint32_t h = hash3(i);
      const int64_t cand = head[h];
      head[h] = (int64_t)i;
      if (cand >= 0 && i - (size_t)cand <= 32768) {
        const size_t lim = n - i < 258 ? n - i : 258;
        unsigned l = 0;
        while (l < lim && src[(size_t)cand + l] == src[i + l])
          ++l;
        if (l >= 3)
          best = l, dist = (unsigned)(i - (size_t)cand);
      }
    }
    if (best) {
      put_match(w, best, dist);
      for (size_t k = 1; k < best && i + k + 3 <= n; ++k) // positions inside the match stay findable
        head[hash3(i + k)] = (int64_t)(i + k);
      i += best;
    } else {
      put_literal_or_length_symbol(w, src[i]);
      ++i;
    }
  }
  put_literal_or_length_symbol(w, 256); // end of block
  w.flush();
  uint32_t a = 1, b = 0; // Adler-32 of the uncompressed bytes
  for (size_t k = 0; k < n;) {
    const size_t stop = n - k < 5552 ? n : k + 5552;
    for (; k < stop; ++k) {
      a += src[k];
      b += a;
    }
    a %= 65521u;
    b %= 65521u;
  }
  const uint32_t adler = (b << 16) | a;
  for (int s = 24; s >= 0; s -= 8)
    out.push_back((uint8_t)(adler >> s));
}

uint32_t
crc32_update(uint32_t crc, const uint8_t* p, size_t n)
{
  static uint32_t table[256];
  static bool ready = false;
  if (!ready) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k)
        c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      table[i] = c;
    }
    ready = true;
  }
  for (size_t i = 0; i < n; ++i)
    crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
  return crc;
}

bool
write_chunk(FILE* f, const char type[4], const uint8_t* data, size_t n)
{
  uint8_t hdr[8] = { (uint8_t)(n >> 24), (uint8_t)(n >> 16), (uint8_t)(n >> 8), (uint8_t)n,
                     (uint8_t)type[0],   (uint8_t)type[1],   (uint8_t)type[2],  (uint8_t)type[3] };
  uint32_t crc = crc32_update(0xFFFFFFFFu, hdr + 4, 4);
  if (n)
    crc = crc32_update(crc, data, n);
  crc ^= 0xFFFFFFFFu;
  const uint8_t tail[4] = { (uint8_t)(crc >> 24), (uint8_t)(crc >> 16), (uint8_t)(crc >> 8), (uint8_t)crc };
  return fwrite(hdr, 1, 8, f) == 8 && (!n || fwrite(data, 1, n, f) == n) && fwrite(tail, 1, 4, f) == 4;
}

} // namespace

extern "C" int
ptc_write_png(const char* path, const void* rgba, uint32_t width, uint32_t height, int keep_alpha)
{
  if (!path || !rgba || !width || !height)
    return PTC_ERR_INVALID;
  const unsigned ch = keep_alpha ? 4u : 3u;
  const size_t row = 1 + (size_t)width * ch;
  if ((uint64_t)row * height > 0x7FFFFFFFull) // one IDAT chunk; 8K RGBA is 133 MB
    return PTC_ERR_INVALID;
  // scanlines, filter type 1 (Sub: each byte minus the same channel of the pixel to its left)
  std::vector<uint8_t> raw(row * height);
  const uint8_t* px = (const uint8_t*)rgba;
  for (uint32_t y = 0; y < height; ++y) {
    uint8_t* d = raw.data() + row * y;
    *d++ = 1;
    uint8_t left[4] = { 0, 0, 0, 0 };
    for (uint32_t x = 0; x < width; ++x) {
      const uint8_t* p = px + ((size_t)y * width + x) * 4;
      for (unsigned c = 0; c < ch; ++c) {
        *d++ = (uint8_t)(p[c] - left[c]);
        left[c] = p[c];
      }
    }
  }
  std::vector<uint8_t> z;
  z.reserve(raw.size() / 2 + 64);
  zlib_compress(raw.data(), raw.size(), z);
  if (z.size() > 0x7FFFFFFFull)
    return PTC_ERR_INVALID;
  FILE* f = fopen(path, "wb");
  if (!f)
    return PTC_ERR_INVALID;
  static const uint8_t sig[8] = { 0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n' };
  const uint8_t ihdr[13] = { (uint8_t)(width >> 24),  (uint8_t)(width >> 16),  (uint8_t)(width >> 8),  (uint8_t)width,
                             (uint8_t)(height >> 24), (uint8_t)(height >> 16), (uint8_t)(height >> 8), (uint8_t)height,
                             8, (uint8_t)(keep_alpha ? 6 : 2), 0, 0, 0 }; // 8 bit, RGB(A), deflate, adaptive, no interlace
  bool ok = fwrite(sig, 1, 8, f) == 8 && write_chunk(f, "IHDR", ihdr, 13) && write_chunk(f, "IDAT", z.data(), z.size()) &&
            write_chunk(f, "IEND", nullptr, 0);
  ok = fclose(f) == 0 && ok;
  return ok ? PTC_OK : PTC_ERR_INVALID;
}
